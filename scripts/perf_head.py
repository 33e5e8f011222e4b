"""Times HEAD (mip-NeRF) render_image, fp32 FFMA path against the tcgen05 bf16 path (CUDA events)."""
import argparse, json, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nerf_b200.head import HeadNeRF
from oracle.g1 import pose_spherical

ap = argparse.ArgumentParser()
ap.add_argument("--hw", type=int, default=400); ap.add_argument("--samples", type=int, default=64); ap.add_argument("--iters", type=int, default=3); ap.add_argument("--prec", default="fp32,bf16")
a = ap.parse_args()
pose = pose_spherical(30.0, -30.0, 4.0311).cuda()
co, cr = pose[None, :3, 3].contiguous(), pose[None, :3, :3].contiguous()
for prec in a.prec.split(","):
    torch.manual_seed(3)
    m = HeadNeRF(precision=prec).cuda()
    run = lambda: m.render_image(co, cr, a.hw, a.hw, 112.0 * a.hw / 100, a.samples)
    run(); torch.cuda.synchronize()
    ts = []
    for i in range(a.iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); img, seg = run(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    ms = float(np.median(ts)); n = a.hw * a.hw
    print(json.dumps({"precision": prec, "hw": a.hw, "samples": a.samples, "ms": ms, "rays_per_s": n / ms * 1e3,
                      "mlp_tflops": n * (a.samples - 1) * 601088 / (ms * 1e-3) / 1e12, "img_mean": img.mean().item()}))
