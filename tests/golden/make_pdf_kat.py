"""Known-answer vectors for the hierarchical (inverse-CDF) sampler -- the algorithm promised at
/root/reference/nerf/model.py:191-193, 503-505 and implemented nowhere in the reference tree; what is pinned here is
the published algorithm (NeRF, section 5.2; the public `sample_pdf` semantics):

    edges  e_i = (t_i + t_{i+1}) / 2,  i = 0 .. S-2           (bins between the midpoints of the coarse samples)
    mass   p_i = (w_{i+1} + 1e-5) / sum_k (w_{k+1} + 1e-5),  i = 0 .. S-3   (interior coarse weights)
    cdf    c_0 = 0, c_{i+1} = c_i + p_i
    u in bin i (c_i <= u < c_{i+1})  ->  t = e_i + (u - c_i) / p_i * (e_{i+1} - e_i);  a bin with p_i < 1e-5 divides by 1;
    u = 1 -> the last edge;  the fine samples are then merged with the coarse ones and sorted.

Every expected value below is derived by HAND from those formulas in closed form (exact rationals evaluated in float64 --
none of it calls oracle/g1.py or the CUDA kernel), so the restatement and the kernel are both checked against
something external to this repository's code.  Writes tests/golden/pdf_kat.npz.

    python tests/golden/make_pdf_kat.py
"""
import os
from fractions import Fraction as F

import numpy as np

E = F(float(np.float32(1e-5)))          # the 1e-5 the fp32 code adds, as the exact rational of its fp32 value


def closed_form(t, w, us):
    """The formulas of the docstring in exact rational arithmetic (no searchsorted, no cumsum library call)."""
    t = [F(float(np.float32(x))) for x in t]; w = [F(float(np.float32(x))) for x in w]
    S = len(t)
    e = [(t[i] + t[i + 1]) / 2 for i in range(S - 1)]
    m = [w[i + 1] + E for i in range(S - 2)]
    tot = sum(m)
    p = [x / tot for x in m]
    c = [F(0)]
    for x in p:
        c.append(c[-1] + x)
    out = []
    for u in us:
        u = F(float(np.float32(u)))
        if u >= c[-1]:
            out.append(e[-1]); continue
        i = max(k for k in range(len(p)) if c[k] <= u)
        denom = p[i] if p[i] >= E else F(1)
        out.append(e[i] + (u - c[i]) / denom * (e[i + 1] - e[i]))
    return np.array([float(x) for x in out])


cases = {}

# A. four equal bins over [0.5, 4.5]: cdf = 0, 1/4, 1/2, 3/4, 1 exactly  =>  t = 0.5 + 4 u, also AT the bin edges; u = 1 -> 4.5
tA, wA = [0, 1, 2, 3, 4, 5], [7, 1, 1, 1, 1, 7]                 # the end weights never enter
uA = [0.0, 0.125, 0.25, 0.5, 0.625, 0.99, 1.0]
fA = closed_form(tA, wA, uA)
assert np.allclose(fA, 0.5 + 4 * np.array(uA), atol=1e-12)
cases["A"] = (tA, wA, uA, fA)

# B. masses 1 : 2 : 1 over the edges 0.5, 1.5, 2.5, 3.5 (cdf 0, ~1/4, ~3/4, 1): by hand, up to the 1e-5 floor,
#    u = 0.1 -> 0.5 + 0.1/0.25 = 0.9;  0.3 -> 1.5 + 0.05/0.5 = 1.6;  0.6 -> 1.5 + 0.35/0.5 = 2.2;  0.9 -> 2.5 + 0.15/0.25 = 3.1
tB, wB, uB = [0, 1, 2, 3, 4], [9, 1, 2, 1, 9], [0.1, 0.3, 0.6, 0.9]
fB = closed_form(tB, wB, uB)
assert np.allclose(fB, [0.9, 1.6, 2.2, 3.1], atol=2e-5)
cases["B"] = (tB, wB, uB, fB)

# C. two EMPTY bins in the middle (masses 1 : 0 : 0 : 1): u = 0.25 -> 0.5 + 0.25/0.5 = 1.0; u = 0.75 -> 3.5 + 0.25/0.5 = 4.0;
#    u just inside the first empty bin (c_1 + p_1/2): its mass 5e-6 is below the 1e-5 floor, the division is by 1 => t = 1.5 + 2.5e-6
tC, wC = [0, 1, 2, 3, 4, 5], [3, 1, 0, 0, 1, 3]
c1 = float((1 + E) / (2 + 4 * E)); p1 = float(E / (2 + 4 * E))
uC = [0.25, 0.75, c1 + 0.5 * p1]
fC = closed_form(tC, wC, uC)
assert np.allclose(fC[:2], [1.0, 4.0], atol=2e-5) and abs(fC[2] - 1.5) < 1e-5
cases["C"] = (tC, wC, uC, fC)

# D. the default draw u = linspace(0, 1, 5) on case A: the five bin edges
uD = list(np.linspace(0.0, 1.0, 5))
fD = closed_form(tA, wA, uD)
assert np.allclose(fD, [0.5, 1.5, 2.5, 3.5, 4.5], atol=1e-12)
cases["D"] = (tA, wA, None, fD)

# E. unevenly spaced (jittered) coarse samples, masses 0.2 : 0.5 : 0.3 over the edges 2.25, 3.0, 3.625, 4.875:
#    u = 0.1 -> 2.25 + 0.5 * 0.75 = 2.625;  0.45 -> 3.0 + 0.5 * 0.625 = 3.3125;  0.85 -> 3.625 + 0.5 * 1.25 = 4.25
tE, wE, uE = [2.0, 2.5, 3.5, 3.75, 6.0], [0.0, 0.2, 0.5, 0.3, 0.0], [0.1, 0.45, 0.85]
fE = closed_form(tE, wE, uE)
assert np.allclose(fE, [2.625, 3.3125, 4.25], atol=1e-4)
cases["E"] = (tE, wE, uE, fE)

out = {"names": np.array(sorted(cases))}
for k, (t, w, u, f) in cases.items():
    out[k + "_t"] = np.array(t, np.float32); out[k + "_w"] = np.array(w, np.float32)
    out[k + "_u"] = np.array([], np.float32) if u is None else np.array(u, np.float32)
    out[k + "_fine"] = f.astype(np.float64)
    out[k + "_merged"] = np.sort(np.concatenate([np.array(t, np.float64), f]))
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "pdf_kat.npz")
np.savez(path, **out)
print("wrote", path, {k: v[3].round(6).tolist() for k, v in cases.items()})
