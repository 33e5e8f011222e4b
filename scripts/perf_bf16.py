"""Step and per-kernel times of the bf16 training step (4096 x 64), the bench's calls without the rest of the bench."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import nerf_pth_state_dict, synthetic_batch
from nerf_b200 import _lib
from nerf_b200.model import NeRF
from nerf_b200.train import Trainer
b, s = 4096, 64
m = NeRF(normalize_position=6.0, precision="bf16").cuda(); m.load_state_dict(nerf_pth_state_dict())
tr = Trainer(m, num_samples=s, density_noise_std=1.0, randomly_sample=True, seed=1)
o, d = (x.cuda() for x in synthetic_batch(b, 0)); tgt = torch.rand(b, 3, device="cuda")
for _ in range(3):
    tr.step(o, d, tgt)
L, P = _lib.lib(), _lib.ptr
z = tr._tc_buffers(b, s, o.device); flat = m.flat_parameters(); g = torch.zeros_like(flat)
lad = m._ladder(2.0, 6.0, s, o.device)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def timed(f, n=10):
    f(); torch.cuda.synchronize()
    tot = 0.0
    for _ in range(n):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); f(); e1.record(); torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / n


fwd = lambda: _lib.check(L.nerf_g1_train_fwd_tc(P(z["packed"]), 3, 3, P(o), P(d), b, s, 2.0, 6.0, P(lad), None, None, 1, 1, 0,
                                                6.0, 1, None, P(z["t"]), P(z["sigma"]), P(z["rgb"]), P(z["ws"]), z["n"], _lib.stream()), "fwd")
bwd = lambda stage: (lambda: _lib.check(L.nerf_g1_train_bwd_tc(P(z["packed"]), P(z["packed_bwd"]), P(flat), 3, 3, P(z["d_sigma"]), P(z["d_rgb"]),
                                                                b, s, P(g), P(z["ws"]), z["n"], stage, _lib.stream()), "bwd"))
t_f, t_d, t_w = timed(fwd), timed(bwd(1)), timed(bwd(3))
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize(); e0.record()
for _ in range(50):
    tr.step(o, d, tgt)
e1.record(); torch.cuda.synchronize()
print(f"step {e0.elapsed_time(e1) / 50:.3f} ms | fwd {t_f:.3f}  dgrad {t_d:.3f}  wgrad {t_w:.3f}")

if os.environ.get("PERF_BF16_MODES"):
    # how the bracket is taken matters at the 5 % level: a 256 MiB write right before the kernel leaves L2 full of dirty lines
    def async_flush(f, n=10):
        evs = []
        for _ in range(n):
            flush.zero_()
            a, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); f(); b_.record(); evs.append((a, b_))
        torch.cuda.synchronize()
        return sum(a.elapsed_time(b_) for a, b_ in evs) / n

    def no_flush(f, n=10):
        evs = []
        for _ in range(n):
            a, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); f(); b_.record(); evs.append((a, b_))
        torch.cuda.synchronize()
        return sum(a.elapsed_time(b_) for a, b_ in evs) / n
    for name, f in (("fwd", fwd), ("dgrad", bwd(1)), ("wgrad", bwd(3))):
        print(f"{name}: flushed+sync {timed(f):.4f}  flushed async {async_flush(f):.4f}  no flush {no_flush(f):.4f}")
