"""Times the conditional (C4/D4) bf16 train step and, via ncu-free CUDA events, its G1 counterpart."""
import os, sys, json
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import synthetic_batch
from nerf_b200.model import NeRF
from nerf_b200.train import Trainer
dev = torch.device("cuda")
o, d = (x.to(dev) for x in synthetic_batch(4096, 0)); tgt = torch.rand(4096, 3, device=dev)
g = torch.Generator().manual_seed(4); scenes = torch.rand(8, 4, generator=g) * 2 - 1
st = scenes[torch.randint(0, 8, (4096,), generator=g)].to(dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
out = {}
for name, kw, sx in (("g1", {}, None), ("c4", {"density_inputs": 7}, st)):
    torch.manual_seed(5)
    m = NeRF(normalize_position=6.0, precision="bf16", **kw).to(dev)
    tr = Trainer(m, num_samples=64, seed=1)
    for i in range(3): tr.step(o, d, tgt, states_x=sx)
    ts = []
    for i in range(20):
        flush.zero_(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); tr.step(o, d, tgt, states_x=sx); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    out[name] = float(np.median(ts))
print(json.dumps(out))
