"""A few fp32-grade (bf16x3) training steps and renders, for ncu (scripts/profile_r02.sh style): 4096 x 64 step, 800 x 800 render."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import nerf_pth_state_dict, synthetic_batch
from nerf_b200.model import NeRF
from nerf_b200.train import Trainer
from oracle.g1 import pose_spherical, TINY_NERF_FOCAL
m = NeRF(normalize_position=6.0, precision="bf16x3").cuda(); m.load_state_dict(nerf_pth_state_dict())
tr = Trainer(m, num_samples=64, seed=1)
o, d = (x.cuda() for x in synthetic_batch(4096, 0)); tgt = torch.rand(4096, 3, device="cuda")
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 3):
    tr.step(o, d, tgt)
pose = pose_spherical(30.0, -30.0, 4.0311).cuda()
with torch.no_grad():
    m.render_image(pose[None, :3, 3].contiguous(), pose[None, :3, :3].contiguous(), 400, 400, TINY_NERF_FOCAL * 4, 2.0, 6.0, 64)
torch.cuda.synchronize()
print("ok")

if os.environ.get("PERF_X3_TIME"):
    # step and per-kernel times (CUDA events over 10 launches each, L2 not flushed: the step's own data exceeds it)
    from nerf_b200 import _lib
    L, P = _lib.lib(), _lib.ptr
    z = tr._tc_buffers(4096, 64, o.device); flat = m.flat_parameters(); g = torch.zeros_like(flat)
    lad = m._ladder(2.0, 6.0, 64, o.device)

    def timed(f, n=10):
        f(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            f()
        e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n
    fwd = lambda: _lib.check(L.nerf_g1_train_fwd_x3(P(z["packed"]), P(o), P(d), 4096, 64, 2.0, 6.0, P(lad), None, None, 1, 1, 0, 6.0, 1,
                                                    P(z["t"]), P(z["sigma"]), P(z["rgb"]), P(z["ws"]), z["n"], _lib.stream()), "fwd")
    bwd = lambda stage: (lambda: _lib.check(L.nerf_g1_train_bwd_x3(P(z["packed"]), P(z["packed_bwd"]), P(flat), P(z["d_sigma"]), P(z["d_rgb"]), 4096, 64,
                                                                    P(g), P(z["ws"]), z["n"], stage, _lib.stream()), "bwd"))
    t_f, t_d, t_dw = timed(fwd), timed(bwd(1)), timed(bwd(2))
    print(f"step {timed(lambda: tr.step(o, d, tgt)):.3f} ms | fwd {t_f:.3f}  dgrad {t_d:.3f}  wgrad x3 + unfold {t_dw - t_d:.3f}")
