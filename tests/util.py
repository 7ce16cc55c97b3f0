"""Shared helpers for the tests (synthetic pairs, error norms)."""
import numpy as np


def std(a):
    a = np.asarray(a, np.float64)
    return (a - a.mean(0)) / a.std(0)


def make_pair(n, seed, kind=0):
    """A standardised synthetic cause-effect pair [n,2] float32, column 0 the cause."""
    rng = np.random.default_rng(seed)
    x = rng.uniform(-1, 1, n) if kind == 0 else rng.normal(size=n)
    if kind == 0:
        y = np.tanh(2.0 * x) + 0.25 * rng.normal(size=n)
    elif kind == 1:
        y = x ** 3 - x + 0.4 * rng.normal(size=n)
    else:
        y = np.abs(x) * rng.normal(size=n) + 0.5 * x
    if n == 1:
        return np.array([[0.3, -0.2]], np.float32)
    return np.stack([std(x), std(y)], 1).astype(np.float32)


def relerr(got, ref):
    """max |got-ref| / max |ref|  (norm-wise relative error)."""
    got = np.asarray(got, np.float64)
    ref = np.asarray(ref, np.float64)
    den = np.max(np.abs(ref))
    return float(np.max(np.abs(got - ref)) / (den if den > 0 else 1.0))
