"""Fits and checks (in float32 arithmetic) the sin/cos polynomials of csrc/physics_f32.cuh on |x| <= 1.6."""
import numpy as np
from numpy.polynomial import chebyshev as Ch, polynomial as Pl

L = np.pi / 2 * 1.02


def fit(f, deg, n=4000):
    k = np.arange(n); z = (np.cos(np.pi * (k + 0.5) / n) + 1) / 2 * L * L
    return Ch.Chebyshev.fit(z, f(z), deg, domain=[0, L * L]).convert(kind=Pl.Polynomial).coef


S = fit(lambda z: (np.sin(np.sqrt(z)) / np.sqrt(z) - 1) / z, 3)
C = fit(lambda z: (np.cos(np.sqrt(z)) - 1) / z, 4)
f32 = np.float32
x = np.linspace(-L, L, 2000001).astype(f32); z = x * x
p = f32(S[-1]) * np.ones_like(x)
for c in S[-2::-1]:
    p = (p * z + f32(c)).astype(f32)
s = ((x * z).astype(f32) * p + x).astype(f32)
q = f32(C[-1]) * np.ones_like(x)
for c in C[-2::-1]:
    q = (q * z + f32(c)).astype(f32)
cc = (q * z + f32(1)).astype(f32)
print("sin coefficients (z^0..z^3 of (sin x / x - 1)/z):", [repr(float(f32(c))) for c in S])
print("cos coefficients (z^0..z^4 of (cos x - 1)/z):   ", [repr(float(f32(c))) for c in C])
print("max abs err  sin %.3e  cos %.3e" % (np.abs(s - np.sin(x.astype(np.float64))).max(), np.abs(cc - np.cos(x.astype(np.float64))).max()))
