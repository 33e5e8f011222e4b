#!/usr/bin/env python
"""fp32 (FFMA) training step and render timing: CUDA events, L2 flushed between steps."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import nerf_pth_state_dict, synthetic_batch, timed_steps  # noqa: E402
from nerf_b200.model import NeRF  # noqa: E402
from nerf_b200.train import Trainer  # noqa: E402

dev = torch.device("cuda")
m = NeRF(normalize_position=6.0, precision="fp32").to(dev); m.load_state_dict(nerf_pth_state_dict())
tr = Trainer(m, num_samples=64, seed=1)
o, d = (x.to(dev) for x in synthetic_batch(4096, 0))
tgt = torch.rand(4096, 3, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for _ in range(3):
    tr.step(o, d, tgt)
ms = timed_steps(lambda i: tr.step(o, d, tgt), 10, flush)
print(f"fp32 train step 4096 x 64: {np.median(ms):.3f} ms  ({4096 / np.median(ms) * 1e3 / 1e3:.1f} k rays/s, "
      f"{262144 * 3_704_832 / np.median(ms) / 1e9:.1f} TFLOP/s)")
with torch.no_grad():
    m.render_rays(o, d, 2.0, 6.0, 64)
    ms = timed_steps(lambda i: m.render_rays(o, d, 2.0, 6.0, 64), 10, flush)
print(f"fp32 render_rays 4096 x 64: {np.median(ms):.3f} ms ({262144 * 1_261_568 / np.median(ms) / 1e9:.1f} TFLOP/s)")
