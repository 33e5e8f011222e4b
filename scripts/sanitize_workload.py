"""Small invocations of every tcgen05 / mbarrier / TMA kernel (and the scan kernels) for compute-sanitizer
(scripts/run_sanitizers.sh): sizes chosen so that each protocol runs several tiles per CTA pair but the whole
script stays short under the sanitizer's slowdown."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from nerf_b200.head import HeadNeRF          # noqa: E402
from nerf_b200.model import NeRF             # noqa: E402
from nerf_b200.train import Trainer          # noqa: E402

z = np.load(os.path.join(ROOT, "tests", "golden", "nerf_pth.npz"))
sd = {str(k): torch.from_numpy(z[str(k)]) for k in z["__keys__"]}
g = torch.Generator().manual_seed(0)
b, s = int(os.environ.get("SAN_RAYS", "40")), 64
o = (torch.randn(b, 3, generator=g) * 0.1 + torch.tensor([0.0, 0.0, 4.0])).cuda()
d = torch.nn.functional.normalize(torch.randn(b, 3, generator=g) * 0.2 + torch.tensor([0.0, 0.0, -1.0]), dim=-1).cuda()
tgt = torch.rand(b, 3, generator=g).cuda()

m = NeRF(normalize_position=6.0, precision="bf16").cuda(); m.load_state_dict(sd)
with torch.no_grad():
    px = m.render_rays(o, d, 2.0, 6.0, s)                                   # g1_fused_tc_kernel<2,false> + composite
    img = m.render_image(o[:1], torch.eye(3, device="cuda")[None], 9, 7, 20.0, 2.0, 6.0, s, num_fine=32)   # hierarchical: pdf merge
tr = Trainer(m, num_samples=s, precision="bf16")
for _ in range(2):
    loss = tr.step(o, d, tgt)                                               # fwd<train>, composite_mse_train, bwd, wgrad, unfold, adam
pix = m.render_rays(o, d, 2.0, 6.0, s, randomly_sample=True, density_noise_std=1.0)   # the autograd node
((pix - tgt) ** 2).mean().backward()
cm = NeRF(normalize_position=6.0, density_inputs=7, precision="bf16").cuda()
ctr = Trainer(cm, num_samples=s, precision="bf16")
ctr.step(o, d, tgt, states_x=torch.rand(b, 4, device="cuda"))               # conditional: ray bias, dz ray sums, state gradients
m32 = NeRF(normalize_position=6.0, precision="fp32").cuda(); m32.load_state_dict(sd)
p32 = m32.render_rays(o[:8], d[:8], 2.0, 6.0, s, randomly_sample=True, density_noise_std=1.0)
((p32 - tgt[:8]) ** 2).mean().backward()                                     # FFMA tiles fwd + bwd
torch.manual_seed(3)
hb = HeadNeRF(precision="bf16").cuda()
with torch.no_grad():
    hb.render_rays(o, d, s)                                                 # head_tc_kernel + head_composite
hf = HeadNeRF().cuda()
hi, hs = hf.render_rays(o[:8], d[:8], 32, randomly_sample=True, density_noise_std=1.0)
(hi.sum() + hs.sum()).backward()                                            # HEAD fp32 fwd + bwd
torch.cuda.synchronize()
print("sanitize workload ok: loss", float(loss), "pixels", float(px.mean()), float(img.mean()))
