#!/usr/bin/env python
"""Fit of the one-MUFU GELU of csrc/umma.cuh (gelu2_f):  GELU(x) = max(x,0) - 0.5|x| 2^-g(|x|),
g(a) ~ -log2 erfc(a / sqrt 2) on [0, 7], degree 6, least squares weighted by the GELU error a deviation of g
causes (0.5 a erfc ln 2) and re-weighted towards the minimax solution.  Prints the coefficients (low to high)
and the error of the fp32 Horner evaluation against the exact GELU.

    python tools/fit_gelu.py
"""
import numpy as np
from scipy.special import erf, erfc

AMAX, DEG = 7.0, 6
a = np.linspace(0, AMAX, 20001)
g = -np.log2(erfc(a / np.sqrt(2)))
w0 = 0.5 * a * erfc(a / np.sqrt(2)) * np.log(2) + 1e-7
V = np.vander(a / AMAX, DEG + 1, increasing=True)
w = w0.copy()
coef = np.linalg.lstsq(V * w[:, None], g * w, rcond=None)[0]
for _ in range(30):
    r = (V @ coef - g) * w
    ww = w * (1 + 5 * np.abs(r) / np.abs(r).max())
    w = ww / ww.max() * w.max()
    coef = np.linalg.lstsq(V * w[:, None], g * w, rcond=None)[0]
c = coef / AMAX ** np.arange(DEG + 1)
x = np.linspace(-9, 9, 400001).astype(np.float32)
ax = np.minimum(np.abs(x), np.float32(AMAX))
p = np.full_like(ax, np.float32(c[-1]))
for k in range(DEG - 1, -1, -1):
    p = (p * ax + np.float32(c[k])).astype(np.float32)
gel = (np.maximum(x, 0) - np.float32(0.5) * ax * np.exp2(-p).astype(np.float32)).astype(np.float32)
ref = 0.5 * x.astype(np.float64) * (1 + erf(x.astype(np.float64) / np.sqrt(2)))
err = np.abs(gel - ref)
print("coefficients c0..c6:", " ".join(f"{v:.8e}" for v in c))
print(f"max |GELU error| = {err.max():.2e} at x = {x[err.argmax()]:.2f}")
