"""Fit the single-MUFU exact-erf GELU used by the tensor-core epilogues.

    gelu(x) = max(x, 0) - |x| * exp2(-h(|x|)),    h(a) = 1 - log2(erfc(a / sqrt 2))
h is smooth on [0, A]; a degree-n polynomial is fitted so that the GELU error (not the error of
h) is minimised in the max norm (Lawson-reweighted least squares in a Chebyshev basis, fp64),
then checked in emulated fp32 Horner arithmetic.  Prints C initialisers.
"""
import numpy as np
from scipy.special import erfc, log_ndtr

A = 6.0


def h_exact(a):
    # Phi(-a) = 0.5 erfc(a / sqrt 2); -log2 of it, computed stably through log_ndtr
    return -log_ndtr(-a) / np.log(2.0)


def fit(deg, n=20001, iters=60):
    a = 0.5 * A * (1 - np.cos(np.pi * (np.arange(n) + 0.5) / n))
    t = 2 * a / A - 1
    V = np.polynomial.chebyshev.chebvander(t, deg)
    y = h_exact(a)
    sens = a * np.exp2(-y) * np.log(2.0)          # d gelu / d h
    w = np.ones(n)
    for _ in range(iters):
        sw = np.sqrt(w) * sens
        c, *_ = np.linalg.lstsq(V * sw[:, None], y * sw, rcond=None)
        e = np.abs((V @ c - y) * sens)
        w = w * (e / e.max() + 1e-3)
        w /= w.sum()
    # convert to monomial coefficients in a
    p = np.polynomial.chebyshev.Chebyshev(c, domain=[0, A]).convert(kind=np.polynomial.Polynomial)
    return p.coef


def gelu_ref(x):
    from scipy.special import ndtr
    return x * ndtr(x)


def check(coef):
    x = np.linspace(-12, 12, 2000001)
    a32 = np.minimum(np.abs(x), A).astype(np.float32)
    c32 = coef.astype(np.float32)
    r = np.full_like(a32, c32[-1])
    for k in c32[-2::-1]:
        r = (r * a32 + k).astype(np.float32)      # fp32 Horner (fma rounding approximated)
    E = np.exp2(-r.astype(np.float64)).astype(np.float32)
    g = np.maximum(x, 0).astype(np.float32) - np.abs(x).astype(np.float32) * E
    err = np.abs(g.astype(np.float64) - gelu_ref(x))
    return err.max(), x[err.argmax()]


if __name__ == "__main__":
    # the kernels use degree 5: positive leading coefficient, so no clamp of the argument is needed
    for deg in (5, 6, 7, 8):
        coef = fit(deg)
        e, at = check(coef)
        print(f"deg {deg}: max |gelu err| = {e:.3e} at x = {at:.3f}")
        print("   ", ", ".join(f"{c:.9e}f" for c in coef))
