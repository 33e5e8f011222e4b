#!/usr/bin/env python
"""Sampler alone: achieved HBM GB/s of nerf_sample_linear (in-kernel Philox: 4S B/ray written; host u: 8S B/ray)."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import load_peaks  # noqa: E402
from nerf_b200 import _lib  # noqa: E402
from scripts.sweep_composite import timed  # noqa: E402

L = _lib.lib(); P = _lib.ptr; peak = load_peaks()["hbm"]
dev = torch.device("cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for lb, s in ((16, 64), (20, 64), (20, 192), (22, 64), (22, 256)):
    b = 1 << lb
    t = torch.empty(b, s, device=dev); uu = torch.rand(b, s, device=dev)
    lad = torch.linspace(2.0, 6.0, s, device=dev)
    for name, fn, by in (("philox", lambda: L.nerf_sample_linear(2.0, 6.0, None, s, b, None, 1, 6, 0, P(t), _lib.stream()), 4),
                         ("philox+ladder", lambda: L.nerf_sample_linear(2.0, 6.0, P(lad), s, b, None, 1, 6, 0, P(t), _lib.stream()), 4),
                         ("host u", lambda: L.nerf_sample_linear(2.0, 6.0, None, s, b, P(uu), 1, 6, 0, P(t), _lib.stream()), 8),
                         ("no jitter", lambda: L.nerf_sample_linear(2.0, 6.0, None, s, b, None, 0, 6, 0, P(t), _lib.stream()), 4)):
        ms = timed(fn, flush)
        print(f"2^{lb} x {s:3d} {name:14s} {ms:8.4f} ms {b * s * by / ms / 1e6:7.0f} GB/s ({b * s * by / ms / 1e6 / peak:.2f})")
