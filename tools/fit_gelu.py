"""Derives the polynomial used by gelu_erf2 (vit_gemm.cu): degree-6 fit of log2(erfc(t)) on [0, 3.95],
weighted by erfc(t) so that the absolute error of erf is minimised, and reports the resulting GELU
error in float32 arithmetic.  python tools/fit_gelu.py"""
import numpy as np
from scipy.special import erf, erfc

TMAX = 3.95
t = np.cos(np.pi * (np.arange(4000) + 0.5) / 4000) * TMAX / 2 + TMAX / 2
coef = np.polynomial.polynomial.polyfit(t, np.log2(erfc(t)), 6, w=erfc(t))
c = [np.float32(v) for v in coef]
print("c0..c6:", ", ".join(f"{float(v):.9e}f" for v in c))
x = np.linspace(-12, 12, 2000001).astype(np.float32)
z = np.minimum(np.abs(x) * np.float32(0.70710678), np.float32(TMAX)).astype(np.float32)
p = c[6] * np.ones_like(z)
for k in range(5, -1, -1):
    p = (p * z + c[k]).astype(np.float32)
e = np.exp2((p - np.float32(1)).astype(np.float64)).astype(np.float32)
g = np.where(x >= 0, x - x * e, x * e)
ref = 0.5 * x.astype(np.float64) * (1 + erf(x.astype(np.float64) / np.sqrt(2)))
print("max abs error of GELU:", np.abs(g - ref).max())
