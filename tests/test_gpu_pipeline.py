"""`train.HostBatchPipeline`: pinned-host batches prefetched on a copy stream under the previous step, losses read back on
a second stream.  The same steps taken directly and through the pipeline see the same batches in the same order: the first
loss (the same forward on identical parameters) is equal to the last bits; later ones agree to the run-to-run noise of the
step itself (the weight-gradient GEMM accumulates its split-K partial sums with red.global.add, whose order is not fixed)."""
import os
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _run(nerf_pth_state_dict, batches, piped):
    from nerf_b200.model import NeRF
    from nerf_b200.train import HostBatchPipeline, Trainer
    m = NeRF(normalize_position=6.0, precision="bf16").cuda(); m.load_state_dict(nerf_pth_state_dict)
    tr = Trainer(m, num_samples=64, density_noise_std=1.0, randomly_sample=True, seed=11, precision="bf16")
    losses = []
    if not piped:
        for hb in batches:
            o, d, t = hb.cuda()
            losses.append(tr.step(o, d, t).item())
    else:
        pipe = HostBatchPipeline(torch.device("cuda"), like=batches[0])
        pipe.submit(batches[0])
        for i in range(len(batches)):
            if i + 1 < len(batches):
                pipe.submit(batches[i + 1])
            o, d, t = pipe.get()
            loss = tr.step(o, d, t)
            pipe.done(loss, ring=tr.loss_ring)
        losses = pipe.losses().tolist()
    torch.cuda.synchronize()
    return m.flat_parameters().clone(), losses


def test_pipeline_steps_equal_direct_steps(nerf_pth_state_dict):
    from bench import synthetic_batch
    from nerf_b200.train import HostBatchPipeline
    batches = []
    for i in range(6):
        o, d = synthetic_batch(512, seed=40 + i)
        tgt = torch.rand(512, 3, generator=torch.Generator().manual_seed(60 + i))
        batches.append(torch.stack([o, d, tgt]).pin_memory())
    p_direct, l_direct = _run(nerf_pth_state_dict, batches, piped=False)
    p_piped, l_piped = _run(nerf_pth_state_dict, batches, piped=True)
    p_again, l_again = _run(nerf_pth_state_dict, batches, piped=False)           # the step's own run-to-run noise
    noise = (p_direct - p_again).abs().max().item()
    diff = (p_direct - p_piped).abs().max().item()
    print(f"max |param| difference: direct vs piped {diff:.2e}; direct vs direct {noise:.2e}; losses {l_piped}")
    assert len(l_piped) == 6 and l_piped[0] == pytest.approx(l_direct[0], rel=1e-6)    # (the loss is an atomicAdd of block sums)
    assert l_piped == pytest.approx(l_direct, rel=1e-3)
    assert diff <= max(4.0 * noise, 6 * 2.5e-4)                                  # six Adam steps of lr 1e-4: a sign flip of a ~zero gradient moves 2e-4
    # the trainer rotates LOSS_RING loss slots and the pipeline waits only for the read-back of the slot about to be
    # overwritten: every loss read back is still the loss of ITS step (12 steps = three trips around the ring)
    from nerf_b200.model import NeRF
    from nerf_b200.train import Trainer
    m = NeRF(normalize_position=6.0, precision="bf16").cuda(); m.load_state_dict(nerf_pth_state_dict)
    tr = Trainer(m, num_samples=64, density_noise_std=0.0, randomly_sample=False, seed=3, precision="bf16")
    assert tr.LOSS_RING == 4
    pipe = HostBatchPipeline(torch.device("cuda"), like=batches[0])
    pipe.submit(batches[0])
    direct = []
    for i in range(12):
        if i + 1 < 12:
            pipe.submit(batches[(i + 1) % 6])
        o, d, t = pipe.get()
        loss = tr.step(o, d, t, apply_update=False)           # no update: the loss depends on the batch only
        assert tr.loss_ring == 4
        pipe.done(loss, ring=tr.loss_ring)
        direct.append(loss)                                    # a view of the slot: read AFTER the loop it shows a later step's loss
    got = pipe.losses().tolist()
    torch.cuda.synchronize()
    assert len(got) == 12
    for i in range(6):
        assert got[i] == pytest.approx(got[i + 6], rel=1e-6), (i, got)       # same batch, same parameters
    assert len({round(x, 7) for x in got[:6]}) == 6                          # and six different batches gave six different losses
    assert direct[0].data_ptr() == direct[4].data_ptr() != direct[1].data_ptr()
    with pytest.raises(ValueError):
        pipe.done(None, ring=9)
    # misuse is an error, not a silent overwrite of a buffer a step still reads
    pipe = HostBatchPipeline(torch.device("cuda"), like=batches[0], depth=2)
    pipe.submit(batches[0]); pipe.submit(batches[1])
    with pytest.raises(RuntimeError):
        pipe.submit(batches[2])
    pipe.get(); pipe.done()
    pipe.submit(batches[2])                                  # a freed buffer is reusable
    with pytest.raises(RuntimeError):
        HostBatchPipeline(torch.device("cuda"), like=batches[0]).get()
