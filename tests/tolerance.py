"""north_star: interpolated features and gradients within 1e-5 RELATIVE error.  For an output that is a sum of signed
terms, "relative to the result" is ill-conditioned where the terms cancel (a gradient of N(0,1) cotangents sums to ~0
for many targets), and a blanket atol hides real errors.  The bound used instead is condition-aware:

    |got - exp| <= tol * sum_k |term_k|          (tol = 1e-5)

with exp the oracle's float64 sum and sum|term| obtained by running the SAME oracle on the absolute values of the
inputs (all of these ops are sums of products of their inputs)."""
import numpy as np


def _np(a):
    return a.detach().cpu().numpy() if hasattr(a, "detach") else np.asarray(a)


def assert_sum_close(got, exp, abs_terms, tol=1e-5, what=""):
    got, exp, bound = _np(got).astype(np.float64), _np(exp).astype(np.float64), _np(abs_terms).astype(np.float64)
    assert got.shape == exp.shape == bound.shape, (what, got.shape, exp.shape, bound.shape)
    err = np.abs(got - exp)
    lim = tol * bound + 1e-30
    bad = err > lim
    if bad.any():
        i = np.unravel_index(np.argmax(err / lim), err.shape)
        raise AssertionError(f"{what}: {bad.sum()} of {bad.size} elements beyond {tol:g} * sum|terms|; worst at {i}: "
                             f"got {got[i]!r} exp {exp[i]!r} err {err[i]:.3e} bound {lim[i]:.3e}")
    return float((err / lim).max())
