#!/usr/bin/env python
"""Per-kernel timing of the bf16 training step (CUDA events, L2 flushed between launches).
    python scripts/perf_train.py [--rays 4096] [--samples 64] [--iters 10]"""
import argparse, json, os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import nerf_pth_state_dict, synthetic_batch, FLOP_PER_POINT_FWD, FLOP_PER_POINT_FWD_BWD  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rays", type=int, default=4096)
    ap.add_argument("--samples", type=int, default=64)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--only", default="")
    a = ap.parse_args()
    from nerf_b200 import _lib, ops
    from nerf_b200.model import NeRF
    from nerf_b200.train import Trainer
    L = _lib.lib(); P = _lib.ptr
    dev = torch.device("cuda")
    m = NeRF(normalize_position=6.0, precision="bf16").to(dev); m.load_state_dict(nerf_pth_state_dict())
    tr = Trainer(m, num_samples=a.samples, seed=1)
    b, s = a.rays, a.samples
    o, d = (x.to(dev) for x in synthetic_batch(b, 0))
    tgt = torch.rand(b, 3, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    z = tr._tc_buffers(b, s, dev)
    flat = m.flat_parameters(); st = _lib.stream
    ladder = m._ladder(2.0, 6.0, s, dev)
    calls = {
        "pack": lambda: (L.nerf_g1_pack_weights(P(flat), 3, 3, P(z["packed"]), st()), L.nerf_g1_pack_weights_bwd(P(flat), 3, 3, P(z["packed_bwd"]), st())),
        "fwd": lambda: L.nerf_g1_train_fwd_tc(P(z["packed"]), 3, 3, P(o), P(d), b, s, 2.0, 6.0, P(ladder), None, None, 1, 1, 0, 6.0, 1, None,
                                               P(z["t"]), P(z["sigma"]), P(z["rgb"]), P(z["ws"]), z["n"], st()),
        "composite_fwd": lambda: L.nerf_composite_fwd(P(z["sigma"]), P(z["rgb"]), None, P(o), P(d), P(z["t"]), b, s, P(z["pixels"]), None, None, None, st()),
        "mse": lambda: L.nerf_mse_loss(P(z["pixels"]), P(tgt), b * 3, P(z["loss"]), P(z["g_pixels"]), st()),
        "composite_bwd": lambda: L.nerf_composite_bwd(P(z["sigma"]), P(z["rgb"]), None, P(o), P(d), P(z["t"]), b, s, P(z["g_pixels"]), None, None, None, P(z["pixels"]), None, None,
                                                      P(z["d_sigma"]), P(z["d_rgb"]), st()),
        "bwd_dgrad": lambda: L.nerf_g1_train_bwd_tc(P(z["packed"]), P(z["packed_bwd"]), P(flat), 3, 3, P(z["d_sigma"]), P(z["d_rgb"]), b, s,
                                                    P(tr.flat_grad), P(z["ws"]), z["n"], 1, st()),
        "bwd_dgrad+wgrad": lambda: L.nerf_g1_train_bwd_tc(P(z["packed"]), P(z["packed_bwd"]), P(flat), 3, 3, P(z["d_sigma"]), P(z["d_rgb"]), b, s,
                                                          P(tr.flat_grad), P(z["ws"]), z["n"], 2, st()),
        "wgrad": lambda: L.nerf_g1_train_bwd_tc(P(z["packed"]), P(z["packed_bwd"]), P(flat), 3, 3, P(z["d_sigma"]), P(z["d_rgb"]), b, s,
                                                P(tr.flat_grad), P(z["ws"]), z["n"], 3, st()),
        "bwd_all": lambda: L.nerf_g1_train_bwd_tc(P(z["packed"]), P(z["packed_bwd"]), P(flat), 3, 3, P(z["d_sigma"]), P(z["d_rgb"]), b, s,
                                                  P(tr.flat_grad), P(z["ws"]), z["n"], 0, st()),
        "step": lambda: tr.step(o, d, tgt),
    }
    out = {}
    for name, fn in calls.items():
        if a.only and name not in a.only.split(","):
            if name in ("pack", "fwd", "composite_fwd", "mse", "composite_bwd", "bwd_dgrad"):
                fn()
            continue
        fn(); fn()
        ts = []
        for i in range(a.iters):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        out[name] = float(np.median(ts))
    pts = b * s
    out["fwd_tflops"] = FLOP_PER_POINT_FWD * pts / (out.get("fwd", np.nan) * 1e-3) / 1e12
    out["step_rays_per_s"] = b / (out.get("step", np.nan) * 1e-3)
    out["mlp_fwd_bwd_tflops"] = FLOP_PER_POINT_FWD_BWD * pts / ((out.get("fwd", np.nan) + out.get("bwd_all", np.nan)) * 1e-3) / 1e12
    print(json.dumps(out))


if __name__ == "__main__":
    main()
