"""A CPU stand-in for cgnn_b200.Engine used ONLY to test host logic (batching, sharding, search
replay) without a GPU: scores are cheap deterministic functions of the instance identity."""
import hashlib

import numpy as np


def _h(*parts):
    d = hashlib.blake2b(repr(parts).encode(), digest_size=8).digest()
    return int.from_bytes(d, "little") / 2.0 ** 64


def edge_cost(a, b):
    return _h("edge", str(a), str(b))


def graph_run_score(key, run):
    """key: sorted tuple of (cause, effect) names."""
    # every present edge costs (its pseudo-random cost - 0.45): some removals / additions / reversals pay off,
    # most do not, and the graph never empties
    return 1.0 + 0.05 * sum(edge_cost(a, b) - 0.45 for a, b in key) + 1e-3 * _h("run", key, run)


def pair_run_score(uid, checksum):
    return 0.02 + 0.05 * _h("pair", int(uid), round(float(checksum), 6))


class FakeEngine(object):
    def __init__(self, device=0):
        self.device = device
        self.calls = []

    def close(self):
        pass

    def device_info(self):
        return dict(n_sm=8, ctas_per_sm_pairwise=2)

    def run_pairwise(self, xy, n_rows=None, n_runs=1, uid_base=0, **kw):
        xy = np.asarray(xy)
        self.calls.append(("pairwise", xy.shape[0], n_runs, int(uid_base)))
        out = np.empty((xy.shape[0], n_runs, 2), np.float32)
        for p in range(xy.shape[0]):
            n = int(n_rows[p]) if n_rows is not None else xy.shape[1]
            for r in range(n_runs):
                for d in range(2):
                    cause = xy[p, :n, d]
                    out[p, r, d] = pair_run_score(uid_base + (p * n_runs + r) * 2 + d,
                                                  float(np.dot(cause, np.arange(n) % 7)))
        return dict(scores=out, trace=None, params=None)

    def run_graph(self, data, plans, n_runs=1, row_subsets=None, uid_base=0, graph_uids=None, **kw):
        self.calls.append(("graph", len(plans), n_runs))
        out = np.empty((len(plans), n_runs), np.float32)
        for g, plan in enumerate(plans):
            cols = plan["columns"]
            key = tuple(sorted((str(cols[p]), str(cols[nd["col"]])) for nd in plan["nodes"] for p in nd["parents"]))
            for r in range(n_runs):
                out[g, r] = graph_run_score(key, r)
        return dict(scores=out, trace=None, params=None)
