"""How far does the EXACT (fp32 FFMA) gradient move when the weights move by a relative eps?  Separates an error of a
reduced-precision backward from the conditioning of the gradient itself (ReLU / LayerNorm-backward cancellation): a
tensor-core forward that is 1e-5 from the fp32 forward cannot give a gradient closer than this to the fp32 one."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import nerf_pth_state_dict, synthetic_batch
from nerf_b200.model import NeRF
from nerf_b200.train import Trainer

sd = nerf_pth_state_dict()
for b, s in ((300, 64), (1024, 64)):
    o, d = (x.cuda() for x in synthetic_batch(b, 5))
    gen = torch.Generator().manual_seed(b + s)
    tgt = torch.rand(b, 3, generator=gen).cuda()
    u = torch.rand(b, s, generator=gen).cuda(); noise = torch.randn(b, s, generator=gen).cuda()

    def grad(prec, eps, seed=0):
        m = NeRF(normalize_position=6.0, precision=prec).cuda(); m.load_state_dict(sd)
        if eps:
            with torch.no_grad():
                f = m.flat_parameters()
                f.mul_(1.0 + eps * torch.randn(f.shape, generator=torch.Generator().manual_seed(seed)).cuda())
        tr = Trainer(m, num_samples=s, precision=prec)
        tr.step(o, d, tgt, u=u, noise=noise, apply_update=False)
        return tr.flat_grad.clone(), m

    g0, m0 = grad("fp32", 0.0)
    names = [k for k, _ in m0.named_parameters()]
    def report(tag, g):
        whole = ((g - g0).norm() / g0.norm()).item()
        per = {k: ((g[o_:o_ + p.numel()] - g0[o_:o_ + p.numel()]).norm() / g0[o_:o_ + p.numel()].norm()).item()
               for (k, p), o_ in zip(m0.named_parameters(), m0._offsets)}
        pick = ["density.weight", "block_2.0.weight", "block_1.11.weight", "block_1.9.weight", "block_1.0.weight", "block_0.0.weight"]
        print(f"{b}x{s} {tag:28s} whole {whole:.2e}  " + "  ".join(f"{k.replace('block_', 'b').replace('.weight', '')} {per[k]:.1e}" for k in pick))
    for eps in (1e-7, 1e-6, 1e-5, 1e-4):
        report(f"fp32, weights x (1 + {eps:g} N)", grad("fp32", eps)[0])
    report("bf16x3", grad("bf16x3", 0.0)[0])
    report("bf16", grad("bf16", 0.0)[0])
