"""Batched evaluation of candidate graphs for the hill-climbing drivers.

The reference's searches (CGNN.py:198-252, CGNN_confounders.py:219-383) are strictly sequential:
one candidate graph = nb_runs instances at a time, so a GPU would idle.  Here the search code is
written as the plain sequential algorithm against a `GraphScorer`, and `run_search` executes it by
MEMOISED RE-EXECUTION:

  1. run the algorithm on a copy of the start graph; `scorer(test_graph)` answers from a cache;
  2. on the first cache miss the run switches to speculation: the miss and every later request of
     the current sweep are recorded and answered NaN - every comparison with NaN is false, so the
     move is rejected (the common outcome) and the 9999 "not evaluated" sentinel of the
     hidden-variable search (CGNN_confounders.py:330,349) cannot win against it either;
     the run is abandoned at the end of that sweep;
  3. all recorded candidates are scored in ONE engine batch (candidates x runs instances), the cache
     is filled, and the algorithm is re-run from the start - host work only, microseconds;
  4. repeat until a run finishes without a miss; that run's graph is the result.

A candidate's score is a deterministic function of (seed, graph structure): its Philox streams are
keyed by a hash of the edge set and its plan is built in canonical (data-column) order, so the same
edge set reached through a different sequence of moves is the same model with the same initial
weights.  So a candidate scored speculatively and then not needed changes
nothing, and the final run takes exactly the decisions a one-by-one evaluation would take:
first-improvement order, tabu list, strict `<` acceptance are all those of the sequential code.
"""
import hashlib
from copy import deepcopy

import numpy as np

from . import runtime
from .plan import build_plan
from .utils.Settings import SETTINGS


class _Speculating(Exception):
    pass


LOG = None      # list of held-back log lines while run_search re-executes an algorithm (see CGNN._say)


def structure_key(graph):
    """Order-free identity of a directed graph: the sorted tuple of its edges."""
    return tuple(sorted((str(a), str(b)) for a, b in graph.get_list_edges(order_by_weight=False, return_weights=False)))


def structure_uid(key):
    """64-bit Philox stream base of a structure; the low 8 bits are left free for the run index."""
    h = hashlib.blake2b(repr(key).encode(), digest_size=8).digest()
    return int.from_bytes(h, "little") & 0x7FFFFFFFFFFFFF00


class GraphScorer(object):
    """score(graph) = mean over the finite run scores (+ complexity * #edges for the hidden-variable
    model), CGNN.py:217 / CGNN_confounders.py:236-237."""

    def __init__(self, df_data, confounders=False, **kwargs):
        self.kwargs = {k: v for k, v in kwargs.items() if k not in ("nb_runs", "nb_jobs", "speculation_horizon")}
        self.nb_runs = int(kwargs.get("nb_runs", SETTINGS.NB_RUNS))
        self.confounders = confounders
        self.df = df_data
        self.columns = None
        self.data = None
        self.cache = {}
        self.pending = None          # list of graphs while speculating
        self.batches = []            # sizes of the engine batches (diagnostics)
        self.kernel_ms = 0.0         # device time of this process's share of the batches
        self.n_scored = 0            # candidates scored so far: the `idx` of the reference's progress lines
        self.horizon = int(kwargs.get("speculation_horizon", 0)) or None

    def _bind(self, graph):
        nodes = set(graph.skeleton.get_list_nodes() if self.confounders else graph.get_list_nodes())
        self.columns = [c for c in self.df.columns if c in nodes]
        if len(self.columns) != len(nodes):
            raise KeyError("graph nodes missing from the data: %s" % sorted(nodes - set(self.columns)))
        self.data = np.ascontiguousarray(self.df[self.columns].values, dtype=np.float32)     # CGNN.py:179-181

    def penalty(self, graph):
        if not self.confounders:
            return 0.0
        return float(SETTINGS.complexity_graph_param) * len(graph.get_list_edges(order_by_weight=False))

    def __call__(self, graph):
        key = structure_key(graph)
        if self.pending is None and key in self.cache:
            return self.cache[key] + self.penalty(graph)
        if self.pending is None:
            self.pending = []
        if key in self.cache:                    # speculating, but this one is already known
            return self.cache[key] + self.penalty(graph)
        if all(structure_key(g) != key for g in self.pending):
            if len(self.pending) >= self._horizon():
                raise _Speculating()     # enough candidates to fill the GPUs: score them before guessing further
            self.pending.append(deepcopy(graph))
        return float("nan")

    def _horizon(self):
        """How many unseen candidates one speculative batch may hold: enough instances for ONE full wave of the graph
        kernel (one CTA per SM) on every GPU of the job.  Everything speculated after an accepted move is wasted, and
        once the GPUs are full a larger batch is no faster per candidate: on the bundled 22-variable example the search
        needs 33 candidates one by one (6.0 s); one wave per batch scores 38 (3.5 s), two waves 52 (4.1 s), three 64
        (5.2 s) - tools/bench_horizon.py.  The decisions are the same for every horizon."""
        if self.horizon is None:
            try:
                slots = sum(e.device_info()["n_sm"] for e in runtime.get_engines()) * runtime.dist_info()[1]
            except Exception:
                slots = 148
            self.horizon = max(1, slots // max(1, self.nb_runs))
        return self.horizon

    def end_of_sweep(self):
        # A sweep that ended with unanswered candidates normally ends the speculative run.  If the batch is still
        # short of a GPU wave it keeps going instead: should the algorithm start another sweep (a move was accepted
        # earlier in this one), its candidates are speculated on top - assuming, as everywhere, that the pending
        # ones get rejected - so that a launch is not spent on one or two stragglers.
        if self.pending is not None and len(self.pending) >= self._horizon():
            raise _Speculating()

    def flush(self):
        graphs, self.pending = self.pending, None
        self.batches.append(len(graphs))
        for g, s in zip(graphs, self.score_batch(graphs)):
            self.cache[structure_key(g)] = s

    def score_batch(self, graphs):
        """Mean finite run score of every graph: one sharded engine submission."""
        if self.data is None:
            self._bind(graphs[0])
        plans = [build_plan(g, columns=self.columns, confounders=self.confounders, canonical=True) for g in graphs]
        keys = [structure_key(g) for g in graphs]
        n_total, cap = self.data.shape[0], int(SETTINGS.max_nb_points)
        runs = self.run_scores(plans, keys, n_total, cap)
        out = []
        for row in runs:
            fin = row[np.isfinite(row)]
            out.append(float(np.mean(fin.astype(np.float64))) if fin.size else float("nan"))
        return out

    def run_scores(self, plans, keys, n_total, cap):
        """[n_graphs, nb_runs] raw run scores.  Graphs are sharded over ranks / GPUs; every graph brings
        its own Philox stream base (a hash of its structure), so its scores do not depend on which
        batch, rank or GPU happened to evaluate it."""
        nb_runs = self.nb_runs
        if nb_runs > 256:
            raise ValueError("nb_runs > 256 is not supported by the structure-keyed random streams")
        uids = [structure_uid(k) for k in keys]

        def call(engine, lo, hi, _uid):
            subsets = None
            if n_total > cap:        # CGNN.py:183-185: a fresh random row subset per (graph, run)
                subsets = np.stack([
                    np.random.default_rng([int(SETTINGS.seed), uids[g] >> 8, r, 0xC6]).permutation(n_total)[:cap]
                    for g in range(lo, hi) for r in range(nb_runs)]).astype(np.int32)
            kw = dict(self.kwargs)
            progress = bool(kw.pop("verbose", SETTINGS.verbose))
            every = 500 if self.confounders else 100            # CGNN.py:123-127, CGNN_confounders.py:142-146
            if progress:
                kw.setdefault("trace_every", every)
            r = engine.run_graph(self.data, plans[lo:hi], n_runs=nb_runs, row_subsets=subsets,
                                 graph_uids=uids[lo:hi], want_trace=progress, **kw)
            self.kernel_ms += float(getattr(engine, "last_kernel_ms", 0.0))
            if progress and r.get("trace") is not None:
                from .GNN import print_progress
                print_progress(r["trace"], int(kw.get("train_epochs", SETTINGS.train_epochs)), every,
                               lambda k: (self.n_scored + lo + k // nb_runs, k % nb_runs))
            return r["scores"]

        out = runtime.run_sharded(len(plans), 0, 0, call, tail_shape=(nb_runs,))
        self.n_scored += len(plans)
        return out


def _run_scores_single(self, plan, uid):
    """One run of one graph on the first local GPU (per-instance API parity)."""
    n_total, cap = self.data.shape[0], int(SETTINGS.max_nb_points)
    subsets = None
    if n_total > cap:
        subsets = np.random.default_rng([int(SETTINGS.seed), uid >> 8, uid & 0xFF, 0xC6]).permutation(n_total)[:cap]
        subsets = subsets.astype(np.int32)[None]
    r = runtime.get_engines()[0].run_graph(self.data, [plan], n_runs=1, row_subsets=subsets, graph_uids=[uid], **self.kwargs)
    return r["scores"][0, 0]


GraphScorer.run_scores_single = _run_scores_single


def run_search(algorithm, graph, scorer, max_rounds=10000):
    """Execute `algorithm(graph_copy, scorer)` by memoised re-execution; returns its final graph."""
    global LOG
    try:
        for _ in range(max_rounds):
            work = deepcopy(graph)
            LOG = []
            try:
                result = algorithm(work, scorer)
                if scorer.pending is None:
                    lines, LOG = LOG, None
                    for line in lines:                 # the one complete execution: every decision printed once
                        print(line)
                    return result
            except _Speculating:
                pass
            LOG = None
            scorer.flush()
    finally:
        LOG = None
    raise RuntimeError("graph search did not converge")
