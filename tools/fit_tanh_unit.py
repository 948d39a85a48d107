"""Refits the odd polynomial tanh(x) ~ x * P(x^2) on [-1, 1] used by csrc/policy.cuh:tanh_unit (Lawson-reweighted least
squares, evaluated in float32 Horner form like the kernel)."""
import numpy as np
x = np.linspace(1e-6, 1.0, 200001); y = np.tanh(x) / x; t = x * x
for deg in (3, 4, 5):
    w = np.ones_like(x)
    for _ in range(60):
        c = np.polynomial.polynomial.polyfit(t, y, deg, w=w)
        err = np.abs(np.polynomial.polynomial.polyval(t, c) * x - np.tanh(x))
        w = w * (1 + 5 * err / err.max()); w /= w.mean()
    c32 = c.astype(np.float32); xf = x.astype(np.float32); tf = xf * xf
    p = np.float32(c32[-1])
    for k in range(deg - 1, -1, -1):
        p = p * tf + c32[k]
    print(deg, "max abs err (float32 Horner):", np.abs((p * xf).astype(np.float64) - np.tanh(xf.astype(np.float64))).max(),
          [float(v) for v in c32])
