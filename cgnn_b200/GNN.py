"""Pairwise Generative Neural Network scorer on the B200 engine.

API of the reference's Code/cgnn/GNN.py:170-207 (`GNN(backend=...)`, `predict_proba(a, b, idx,
**kwargs)`) plus the per-instance function `b200_run_instance` with the signature of
`tf_run_instance` (GNN.py:135-166).  The reference runs `nb_runs` processes per pair, each
building two TF graphs and calling `sess.run` 3000 times; here all (pair, run, direction) instances
of a call go to the GPU in one launch (`cgnn_b200_run_pairwise`).

kwargs override SETTINGS exactly as in the reference: nb_runs, h_layer_dim, train_epochs,
test_epochs, learning_rate, init_std (GNN.py:45-48,90,112,187-188); `max_nb_points` is read from
SETTINGS only (GNN.py:150).  `nb_jobs` is accepted and ignored (no process pool).
"""
import numpy as np

from . import runtime
from .PairwiseModel import Pairwise_Model
from .utils.Settings import SETTINGS

_BACKENDS = ("B200", "TensorFlow", "PyTorch")     # the reference's names are aliases of the only engine


def _subsample_runs(m, n_runs, uid, run0=0):
    """GNN.py:150-153: a pair longer than SETTINGS.max_nb_points is cut to that many random rows, once
    per run (both directions of a run share the rows).  The rows are a function of (seed, pair index, run), so
    the per-instance seam and the batched path draw the same ones.  Returns [n_runs, max, 2]."""
    cap = int(SETTINGS.max_nb_points)
    out = np.empty((n_runs, cap, 2), np.float32)
    for r in range(n_runs):
        rng = np.random.default_rng([int(SETTINGS.seed), int(uid), run0 + r, 0x5AB5])
        out[r] = m[rng.permutation(m.shape[0])[:cap]]
    return out


def _staging(shape):
    """Zeroed float32 host batch buffer; page-locked when a CUDA runtime is present so that the engine's
    H2D copy is a true asynchronous DMA."""
    try:
        import torch
        if torch.cuda.is_available():
            return torch.zeros(shape, dtype=torch.float32).pin_memory().numpy()
    except Exception:
        pass
    return np.zeros(shape, np.float32)


def print_progress(trace, train_epochs, every, ident):
    """The reference's progress lines (GNN.py:98-102,119-121; CGNN.py:123-127): one per instance and `every`-th
    iteration of training (the pre-update loss) and of evaluation.  trace: [n_inst, train+test]; ident(k) -> (idx, run)."""
    for k in range(trace.shape[0]):
        idx, run = ident(k)
        for part, lo, hi in (("train", 0, train_epochs), ("test", train_epochs, trace.shape[1])):
            for it in range(0, hi - lo, every):
                print("Pair:{}, Run:{}, Iter:{}, score:{}".format(idx, run, it, trace[k, lo + it]))


def score_pairs(pairs, idx0=0, **kwargs):
    """Per-run direction scores of a list of [N,2] float32 matrices: array [n_pairs, nb_runs, 2]
    (A->B, B->A).  The batched replacement of `[tf_run_instance(m, idx, run) for run in range(nb_runs)]`
    for every pair (GNN.py:193-194)."""
    nb_runs = int(kwargs.get("nb_runs", SETTINGS.NB_RUNS))
    cap = int(SETTINGS.max_nb_points)
    sub_idx0, sub_run0 = kwargs.pop("_subsample_key", (idx0, 0))      # (pair index, first run) of the row-subset streams
    mats = [np.ascontiguousarray(m, np.float32) for m in pairs]
    long_pairs = any(m.shape[0] > cap for m in mats)
    if long_pairs:      # one engine "pair" per (pair, run) because every run owns its row subset
        slots = []
        for k, m in enumerate(mats):
            slots += list(_subsample_runs(m, nb_runs, sub_idx0 + k, sub_run0)) if m.shape[0] > cap else [m] * nb_runs
        runs_per_slot = 1
    else:
        slots, runs_per_slot = mats, nb_runs
    ld = max(m.shape[0] for m in slots)
    xy = _staging((len(slots), ld, 2))
    n_rows = np.empty(len(slots), np.int32)
    for k, m in enumerate(slots):
        xy[k, :m.shape[0]] = m
        n_rows[k] = m.shape[0]
    kw = {k: v for k, v in kwargs.items() if k not in ("nb_runs", "nb_jobs", "n_runs", "verbose")}
    progress = bool(kwargs.get("verbose", SETTINGS.verbose))      # GNN.py:98-102,119-121: a line every 100 iterations
    if progress:
        kw.setdefault("trace_every", 100)

    def call(engine, lo, hi, uid):
        r = engine.run_pairwise(xy[lo:hi], n_rows=n_rows[lo:hi], n_runs=runs_per_slot, uid_base=uid, want_trace=progress, **kw)
        if progress and r.get("trace") is not None:

            def ident(k):                       # trace row k of this call -> (pair index, run) as the reference numbers them
                slot = lo + k // (2 * runs_per_slot)
                if runs_per_slot > 1:
                    return idx0 + slot, (k // 2) % runs_per_slot
                return idx0 + slot // nb_runs, slot % nb_runs

            print_progress(r["trace"], int(kw.get("train_epochs", SETTINGS.train_epochs)), 100, ident)
        return r["scores"]

    uid0 = int(kwargs.get("uid_base", 0)) + idx0 * nb_runs * 2
    kw.pop("uid_base", None)
    scores = runtime.run_batch(len(slots), runs_per_slot * 2, uid0, call, tail_shape=(runs_per_slot, 2),
                               costs=[float(n) * float(n) for n in n_rows])
    return scores.reshape(len(mats), nb_runs, 2)


def b200_run_instance(m, idx, run, **kwargs):
    """Drop-in for `tf_run_instance(m, idx, run, **kwargs)` (GNN.py:135-166): `[score A->B, score B->A]`
    of ONE run of one pair.  Provided for API parity; prefer the batched calls."""
    nb_runs = int(kwargs.get("nb_runs", SETTINGS.NB_RUNS))
    kw = dict(kwargs, nb_runs=1, uid_base=int(kwargs.get("uid_base", 0)) + (idx * nb_runs + run) * 2,
              _subsample_key=(idx, run))            # a fresh row subset per (pair, run), as GNN.py:150-153
    s = score_pairs([m], 0, **kw)
    return [float(s[0, 0, 0]), float(s[0, 0, 1])]


def causal_score(run_scores, verbose=False):
    """GNN.py:196-207: (mean BA - mean AB) / (mean BA + mean AB) over the runs of one pair."""
    ab = np.mean(run_scores[:, 0].astype(np.float64))
    ba = np.mean(run_scores[:, 1].astype(np.float64))
    if verbose:
        for s in run_scores[:, 0]:
            print(s)
        print(ab)
        for s in run_scores[:, 1]:
            print(s)
        print(ba)
    return (ba - ab) / (ba + ab)


class GNN(Pairwise_Model):
    """Shallow generative networks fitted in both directions with an MMD loss; the better fit wins."""

    def __init__(self, backend="B200"):
        super(GNN, self).__init__()
        if backend not in _BACKENDS:
            raise KeyError(backend)            # the reference fails the same way, at predict time (GNN.py:182,193)
        self.backend = backend

    def predict_proba(self, a, b, idx=0, **kwargs):
        return self.predict_proba_batch([(a, b)], idx, **kwargs)[0]

    def predict_proba_batch(self, pairs, idx0=0, **kwargs):
        mats = []
        for a, b in pairs:
            a, b = np.asarray(a), np.asarray(b)
            mats.append(np.hstack((a.reshape(-1, 1), b.reshape(-1, 1))).astype("float32"))     # GNN.py:184-190
        verbose = kwargs.pop("verbose", SETTINGS.verbose)
        scores = score_pairs(mats, idx0, **kwargs)
        return [float(causal_score(scores[k], verbose)) for k in range(len(mats))]
