#!/usr/bin/env python
"""SURVEY D5 / BASELINE config 5: compositing and sampling kernel sweep, achieved HBM GB/s against the
measured peak.  B in 2^12..2^22 rays, S in {32,64,128,256}; inputs sigma = 10 relu(N(0,1)), rgb ~ N(0,1),
t stratified in [2,6] (seed 6).  L2 (126 MB) is flushed between timed launches.
    python scripts/sweep_composite.py [--max-log2 22] > profiles/rNN_sweep.json"""
import argparse, json, os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import load_peaks  # noqa: E402


def timed(fn, flush, iters=5):
    fn(); fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--max-log2", type=int, default=22)
    a = ap.parse_args()
    from nerf_b200 import _lib
    L = _lib.lib(); P = _lib.ptr
    peak = load_peaks()["hbm"]
    dev = torch.device("cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    gen = torch.Generator(device=dev).manual_seed(6)
    rows = []
    for lb in range(12, a.max_log2 + 1, 2):
        for s in (32, 64, 128, 256):
            b = 1 << lb
            if b * s * 40 > 60e9:
                continue
            sigma = torch.randn(b, s, device=dev, generator=gen).relu_() * 10
            rgb = torch.randn(b, s, 3, device=dev, generator=gen)
            o = torch.zeros(b, 3, device=dev); d = torch.randn(b, 3, device=dev, generator=gen)
            t = torch.empty(b, s, device=dev)
            st = _lib.stream
            f_s = lambda: L.nerf_sample_linear(2.0, 6.0, None, s, b, None, 1, 6, 0, P(t), st())
            ms_s = timed(f_s, flush)
            uu = torch.rand(b, s, device=dev, generator=gen)
            f_su = lambda: L.nerf_sample_linear(2.0, 6.0, None, s, b, P(uu), 1, 6, 0, P(t), st())
            ms_su = timed(f_su, flush)
            f_s()
            pix = torch.empty(b, 3, device=dev); gp = torch.randn(b, 3, device=dev, generator=gen)
            ds = torch.empty_like(sigma); dc = torch.empty_like(rgb)
            f_f = lambda: L.nerf_composite_fwd(P(sigma), P(rgb), None, P(o), P(d), P(t), b, s, P(pix), None, None, None, st())
            f_b = lambda: L.nerf_composite_bwd(P(sigma), P(rgb), None, P(o), P(d), P(t), b, s, P(gp), None, None, None,
                                               P(pix), None, None, P(ds), P(dc), st())
            ms_f, ms_b = timed(f_f, flush), timed(f_b, flush)
            by_s, by_f, by_b = b * 4 * s, b * (20 * s + 12 + 24), b * (36 * s + 24 + 24)
            rows.append({"rays": b, "samples": s,
                         "sample_ms": ms_s, "sample_gbs": by_s / ms_s / 1e6, "sample_frac": by_s / ms_s / 1e6 / peak,
                         "sample_u_ms": ms_su, "sample_u_gbs": 2 * by_s / ms_su / 1e6, "sample_u_frac": 2 * by_s / ms_su / 1e6 / peak,
                         "fwd_ms": ms_f, "fwd_gbs": by_f / ms_f / 1e6, "fwd_frac": by_f / ms_f / 1e6 / peak,
                         "bwd_ms": ms_b, "bwd_gbs": by_b / ms_b / 1e6, "bwd_frac": by_b / ms_b / 1e6 / peak})
            r = rows[-1]
            print(f"2^{lb} x {s:3d}: sample(philox) {r['sample_gbs']:6.0f} GB/s ({r['sample_frac']:.2f}) sample(u) {r['sample_u_gbs']:6.0f} ({r['sample_u_frac']:.2f})  fwd {r['fwd_gbs']:7.0f} ({r['fwd_frac']:.2f})"
                  f"  bwd {r['bwd_gbs']:7.0f} ({r['bwd_frac']:.2f})", file=sys.stderr)
            del sigma, rgb, t, ds, dc
    print(json.dumps({"peak_hbm_gbs": peak, "bytes": "sample (in-kernel Philox) 4S, sample (host u) 8S; fwd 20S+36; bwd 36S+48 per ray", "rows": rows}))


if __name__ == "__main__":
    main()
