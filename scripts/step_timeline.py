"""The bf16 training step call by call: CUDA events between the C-ABI calls of one step (each event costs ~4 us itself), then
the same sequence without them.  Shows what the launches around the three tensor-core kernels cost."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import nerf_pth_state_dict, synthetic_batch
from nerf_b200 import _lib, ops
from nerf_b200.model import NeRF
from nerf_b200.train import Trainer
b, s = 4096, 64
m = NeRF(normalize_position=6.0, precision="bf16").cuda(); m.load_state_dict(nerf_pth_state_dict())
tr = Trainer(m, num_samples=s, density_noise_std=1.0, randomly_sample=True, seed=1)
o, d = (x.cuda() for x in synthetic_batch(b, 0)); tgt = torch.rand(b, 3, device="cuda")
for _ in range(5):
    tr.step(o, d, tgt)
L, P, st = _lib.lib(), _lib.ptr, _lib.stream
z = tr._tc_buffers(b, s, o.device); flat = m.flat_parameters(); lad = m._ladder(2.0, 6.0, s, o.device)


def bwd(stage):
    return lambda: L.nerf_g1_train_bwd_tc(P(z["packed"]), P(z["packed_bwd"]), P(flat), 3, 3, P(z["d_sigma"]), P(z["d_rgb"]), b, s,
                                          P(tr.flat_grad), P(z["ws"]), z["n"], stage, st())


calls = [
    ("pack_weights (vectors, stages)", lambda: L.nerf_g1_pack_weights(P(flat), 3, 3, P(z["packed"]), st())),
    ("pack_weights_bwd", lambda: L.nerf_g1_pack_weights_bwd(P(flat), 3, 3, P(z["packed_bwd"]), st())),
    ("train_fwd_tc", lambda: L.nerf_g1_train_fwd_tc(P(z["packed"]), 3, 3, P(o), P(d), b, s, 2.0, 6.0, P(lad), None, None, 1, 1, 0, 6.0, 1, None,
                                                    P(z["t"]), P(z["sigma"]), P(z["rgb"]), P(z["ws"]), z["n"], st())),
    ("composite_mse_train", lambda: L.nerf_composite_mse_train(P(z["sigma"]), P(z["rgb"]), P(o), P(d), P(z["t"]), P(tgt), b, s, 1.0, 1, 0,
                                                               P(z["pixels"]), P(z["loss"]), P(z["d_sigma"]), P(z["d_rgb"]), st())),
    ("train_bwd_tc (memset, dgrad, wgrad, unfold)", bwd(0)),
    ("adam", lambda: ops.adam_step(flat, tr.flat_grad, tr.exp_avg, tr.exp_avg_sq, 10, 1e-4)),
]


def seq():
    for _, f in calls:
        f()


for _ in range(3):
    seq()
torch.cuda.synchronize()
N = 20
tot = [0.0] * len(calls); whole = 0.0
for _ in range(N):
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(len(calls) + 1)]
    ev[0].record()
    for i, (_, f) in enumerate(calls):
        f(); ev[i + 1].record()
    torch.cuda.synchronize()
    for i in range(len(calls)):
        tot[i] += ev[i].elapsed_time(ev[i + 1])
    whole += ev[0].elapsed_time(ev[-1])
for (n, _), t in zip(calls, tot):
    print(f"{n:45s} {t / N * 1e3:8.1f} us")
print(f"{'whole step, events between the calls':45s} {whole / N * 1e3:8.1f} us")


def region(f):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f(); torch.cuda.synchronize(); e0.record()
    for _ in range(N):
        f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / N * 1e3


print(f"{'the same sequence without events':45s} {region(seq):8.1f} us")
t2, t0 = region(bwd(2)), region(bwd(0))
print(f"{'unfold_kernel (bwd stage 0 - stage 2)':45s} {t0 - t2:8.1f} us")
