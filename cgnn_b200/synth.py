"""Synthetic inputs for the benchmark configurations of BASELINE.json (configs[3], configs[4]).

Recipe of SURVEY.md section 8(d), itself taken from the reference's (broken, unused) generator
package: uniform cause (Code/cgnn/generators/functions_default.py:24), cubic-spline mechanism
through 4 random knots on [min-std, max+std] (:33-35), additive Gaussian noise of random scale
(:28-29, without the +-2 offset), standardised columns, coin-flip orientation
(random_graph_generator.py:164-175).  Deterministic per pair id.
"""
import numpy as np


def _standardise(v):
    v = np.asarray(v, np.float64)
    s = v.std()
    return (v - v.mean()) / (s if s > 0 else 1.0)


def _spline_mechanism(x, rng, knots=4):
    from scipy.interpolate import UnivariateSpline
    g = np.linspace(x.min() - x.std(), x.max() + x.std(), knots)
    return UnivariateSpline(g, rng.normal(size=knots), k=min(3, knots - 1))(x)


def make_pair(pair_id, n=1500):
    """Returns (xy [n,2] float32 with column 0 = A, column 1 = B, label) with label = +1 if A causes B."""
    rng = np.random.default_rng(20260000 + int(pair_id))
    x = rng.uniform(-1.0, 1.0, n)
    y = _standardise(_spline_mechanism(x, rng)) + 0.7 * rng.uniform() * rng.normal(size=n)
    a, b = _standardise(x), _standardise(y)
    label = 1.0
    if rng.integers(0, 2):
        a, b, label = b, a, -1.0
    return np.stack([a, b], 1).astype(np.float32), label


def make_pairs(first_id, count, n=1500):
    xy = np.empty((count, n, 2), np.float32)
    labels = np.empty(count, np.float32)
    for k in range(count):
        xy[k], labels[k] = make_pair(first_id + k, n)
    return xy, labels


def make_dag_dataset(n_nodes=20, n_rows=2000, max_parents=3, seed=20260500):
    """Layered random DAG (random_graph_generator.py:63-104 recipe): returns (data [n_rows, n_nodes]
    float64 standardised, edges list of (parent, child) index pairs)."""
    rng = np.random.default_rng(seed)
    data = np.zeros((n_rows, n_nodes))
    edges = []
    n_roots = max(2, n_nodes // 5)
    for v in range(n_nodes):
        if v < n_roots:
            data[:, v] = _standardise(rng.uniform(-1, 1, n_rows))
            continue
        k = int(rng.integers(1, max_parents + 1))
        pa = rng.choice(v, size=min(k, v), replace=False)
        acc = np.zeros(n_rows)
        for p in pa:
            acc += _spline_mechanism(data[:, p], rng)
            edges.append((int(p), v))
        data[:, v] = _standardise(_standardise(acc) + 0.4 * rng.uniform(0.3, 1.0) * rng.normal(size=n_rows))
    return data, edges
