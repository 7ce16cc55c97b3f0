"""Causal Generative Neural Network over a whole graph, on the B200 engine.

Keeps the reference's entry points (Code/cgnn/CGNN.py): `run_CGNN_*` per-instance scorer
(:163-195), `hill_climbing(graph, data, run_cgnn_function, **kwargs)` (:198-252) and the `CGNN`
model class (:330-381).  One generative MLP per node, fed by its generated parents and its own
noise, the joint sample compared with the data under the 7-bandwidth MMD; hill climbing reverses
one edge at a time and keeps a reversal iff the mean-over-runs MMD drops.

The search is the same first-improvement sequence of decisions; it is *executed* differently:
candidates are scored in batches by memoised re-execution (see cgnn_b200/search.py), because a
32-instance call cannot fill a B200.
"""
import warnings
from copy import deepcopy

import numpy as np
from pandas import DataFrame

from .GNN import GNN
from .GraphModel import GraphModel
from .search import GraphScorer, run_search, structure_key, structure_uid
from .utils.scaling import scale
from .utils.Settings import SETTINGS

_BACKENDS = ("B200", "TensorFlow", "PyTorch")


def _say(*a):
    """The reference's progress prints.  Inside a batched search the algorithm is re-executed once per batch; its lines
    are held back (search.LOG) and only the final, complete execution is printed - each decision once, as the
    reference prints it."""
    if SETTINGS.verbose:
        from . import search
        if search.LOG is not None:
            search.LOG.append(" ".join(str(x) for x in a))
        else:
            print(*a)


def run_CGNN_b200(df_data, graph, idx=0, run=0, **kwargs):
    """Drop-in for `run_CGNN_tf(df_data, graph, idx, run, **kwargs)` (CGNN.py:163-195): trains and
    evaluates ONE run of one graph and returns its mean test MMD.  Prefer `hill_climbing`, which
    batches; this exists so that code written against the reference's seam keeps working."""
    scorer = GraphScorer(df_data, confounders=False, **dict(kwargs, nb_runs=1))
    scorer._bind(graph)
    from .plan import build_plan
    plan = build_plan(graph, columns=scorer.columns, canonical=True)
    uid = structure_uid(structure_key(graph)) + int(run)
    runs = scorer.run_scores_single(plan, uid)
    return float(runs)


class _FunctionScorer(object):
    """Adapter for a caller-supplied per-instance function with the reference's signature
    `fn(data, graph, idx, run, **kwargs) -> float`: no batching, no speculation."""
    pending = None

    def __init__(self, df_data, fn, penalty=False, **kwargs):
        self.df, self.fn, self.kwargs, self.use_penalty = df_data, fn, kwargs, penalty
        self.nb_runs = int(kwargs.get("nb_runs", SETTINGS.NB_RUNS))
        self.calls = 0

    def __call__(self, graph):
        vals = [self.fn(self.df, graph, self.calls, run, **self.kwargs) for run in range(self.nb_runs)]
        self.calls += 1
        s = np.mean([v for v in vals if np.isfinite(v)])
        if self.use_penalty:
            s += SETTINGS.complexity_graph_param * len(graph.get_list_edges(order_by_weight=False))
        return s

    def end_of_sweep(self):
        pass


def _climb(graph, scorer):
    """The sequential algorithm of CGNN.py:198-252 against an abstract scorer."""
    visited = [graph.get_dict_nw()]
    best = scorer(graph)
    _say("Graph score : " + str(best))
    moved = True
    while moved:
        moved = False
        for edge in graph.get_list_edges():                     # ascending weight, fixed for the sweep
            cause, effect = edge[0], edge[1]
            trial = deepcopy(graph)
            trial.reverse_edge(cause, effect)
            shape = trial.get_dict_nw()
            if trial.is_cyclic() or shape in visited:
                _say("No Evaluation for {}".format([edge]))
                continue
            _say("Edge {} in evaluation :".format(edge))
            visited.append(shape)
            score = scorer(trial)
            _say("Current score : " + str(score))
            _say("Best score : " + str(best))
            if score < best:
                graph.reverse_edge(cause, effect)
                moved = True
                best = score
                _say("Edge {} got reversed !".format(edge))
        scorer.end_of_sweep()
    return graph


def hill_climbing(graph, data, run_cgnn_function=run_CGNN_b200, **kwargs):
    """Greedy first-improvement edge reversal (CGNN.py:198-252).  `graph` is improved in place and
    returned.  With the default scorer the candidates of a sweep are evaluated in batches; a custom
    `run_cgnn_function(data, graph, idx, run, **kwargs) -> float` is called run by run, as the reference does."""
    if run_cgnn_function is run_CGNN_b200:
        scorer = GraphScorer(data, confounders=False, **kwargs)
        final = run_search(_climb, graph, scorer)
        graph._graph = final._graph                              # same object mutated, like the reference
        hill_climbing.last_batches = list(scorer.batches)
        hill_climbing.last_kernel_seconds = scorer.kernel_ms * 1e-3
        return graph
    return _climb(graph, _FunctionScorer(data, run_cgnn_function, **kwargs))


def exploratory_hill_climbing(graph, data, run_cgnn_function=run_CGNN_b200, **kwargs):
    raise NotImplementedError("exploratory_hill_climbing is unrunnable in the reference (CGNN.py:291-292) and out of scope")


def tabu_search(graph, data, run_cgnn_function=run_CGNN_b200, **kwargs):
    raise ValueError("Not Yet Implemented")                      # CGNN.py:327


class CGNN(GraphModel):
    """Generate the whole causal graph with generative networks and improve the edge directions."""

    def __init__(self, backend="B200"):
        super(CGNN, self).__init__()
        if backend not in _BACKENDS:
            print("No backend known as {}".format(backend))      # CGNN.py:348-350
            raise ValueError
        self.backend = backend
        self.infer_graph = run_CGNN_b200

    def create_graph_from_data(self, data):
        print("The CGNN model is not able (yet?) to model the graph directly from raw data")
        raise ValueError

    def orient_directed_graph(self, data, dag, alg="HC", **kwargs):
        """Improve a DAG with CGNN; the data is re-standardised first (CGNN.py:356-367)."""
        data = DataFrame(scale(data.values), columns=data.columns)
        algorithms = {"HC": hill_climbing, "tabu": tabu_search, "EHC": exploratory_hill_climbing}
        return algorithms[alg](dag, data, self.infer_graph, **kwargs)

    def generate(self, data, dag, n_samples=None, run=0, **kwargs):
        """Fit the generative model of `dag` to `data` (one run) and draw a sample from it: the
        `CGNN_tf.generate` of the reference (CGNN.py:155-160), which its model class never exposed.
        Returns a DataFrame [n_samples, n_var] in the (standardised) units of the data."""
        from . import runtime
        from .plan import build_plan, n_params
        data = DataFrame(scale(data.values), columns=data.columns)
        cols = [c for c in data.columns if c in set(dag.get_list_nodes())]
        plan = build_plan(dag, columns=cols, canonical=True)
        X = np.ascontiguousarray(data[cols].values, np.float32)
        eng = runtime.get_engines()[0]
        uid = structure_uid(structure_key(dag)) + int(run)
        kw = {k: v for k, v in kwargs.items() if k not in ("nb_runs", "nb_jobs")}
        fit = eng.run_graph(X, [plan], n_runs=1, graph_uids=[uid], want_params=True, **kw)
        h = int(kwargs.get("h_layer_dim", SETTINGS.h_layer_dim))
        n = int(n_samples or X.shape[0])
        noise = np.random.default_rng([int(SETTINGS.seed), uid >> 8, 0x6E]).standard_normal((plan["n_noise"], n))
        sample = eng.graph_generate(plan, fit["params"][0, :n_params(plan, h)], noise, **kw)
        return DataFrame(sample, columns=cols)

    def orient_undirected_graph(self, data, umg, **kwargs):
        """Orient a skeleton with the pairwise GNN, then improve it with CGNN (CGNN.py:369-381)."""
        warnings.warn("The pairwise GNN model is computed on each edge of the UMG "
                      "to initialize the model and start CGNN with a DAG")
        dag = GNN(backend=self.backend).orient_graph(data, umg, **kwargs)
        return self.orient_directed_graph(data, dag, **kwargs)
