"""Times the tcgen05 fused render (800x800 rays x S samples) with CUDA events; prints rays/s and
MLP TFLOP/s (algorithmic, 1 261 568 FLOP per point)."""
import argparse, json, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nerf_b200 import ops, _lib
from nerf_b200.model import NeRF
from oracle.g1 import pose_spherical

ap = argparse.ArgumentParser()
ap.add_argument("--hw", type=int, default=800); ap.add_argument("--samples", type=int, default=64)
ap.add_argument("--iters", type=int, default=5); ap.add_argument("--fine", type=int, default=0)
a = ap.parse_args()
z = np.load(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "nerf_pth.npz"))
sd = {str(k): torch.from_numpy(z[str(k)]) for k in z["__keys__"]}
m = NeRF(normalize_position=6.0, precision="bf16").cuda(); m.load_state_dict(sd)
pose = pose_spherical(30.0, -30.0, 4.0311).cuda()
co, cr = pose[None, :3, 3].contiguous(), pose[None, :3, :3].contiguous()
focal = 0.5 * a.hw / np.tan(0.5 * 0.6911112)
packed = ops.g1_pack_weights(m.flat_parameters())
n = a.hw * a.hw
def run():
    with torch.no_grad():
        return m.render_image(co, cr, a.hw, a.hw, focal, 2.0, 6.0, a.samples, num_fine=a.fine)
run(); torch.cuda.synchronize()
ts = []
for i in range(a.iters):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); img = run(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
ms = float(np.median(ts))
evals = a.samples + ((a.samples + a.fine) if a.fine else 0)
print(json.dumps({"hw": a.hw, "samples": a.samples, "fine": a.fine, "ms": ms, "rays_per_s": n / ms * 1e3,
                  "mlp_tflops": n * evals * 1261568 / (ms * 1e-3) / 1e12, "all_ms": ts,
                  "img_mean": img.mean().item()}))
