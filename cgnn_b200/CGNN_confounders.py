"""CGNN with hidden common causes (correlated noise), on the B200 engine.

Keeps the reference's entry points (Code/cgnn/CGNN_confounders.py): per-instance scorer (:182-216),
`hill_climbing_confounders` (:219-383) and the `CGNN_confounders` class (:461-512).  Every skeleton
edge carries one noise column shared by both endpoints; the search may reverse, remove or re-add
edges and every score pays `SETTINGS.complexity_graph_param` per edge.
"""
from copy import deepcopy

import numpy as np
from pandas import DataFrame

from .CGNN import _FunctionScorer, _say, _BACKENDS
from .GraphModel import GraphModel
from .search import GraphScorer, run_search, structure_key, structure_uid
from .utils.scaling import scale
from .utils.Settings import SETTINGS

NO_SCORE = 9999        # sentinel of CGNN_confounders.py:330,349


def run_CGNN_confounders_b200(df_data, graph, idx=0, run=0, **kwargs):
    """Drop-in for `run_CGNN_confounders_tf` (CGNN_confounders.py:182-216): one run of one graph."""
    scorer = GraphScorer(df_data, confounders=True, **dict(kwargs, nb_runs=1))
    scorer._bind(graph)
    from .plan import build_plan
    plan = build_plan(graph, columns=scorer.columns, confounders=True, canonical=True)
    return float(scorer.run_scores_single(plan, structure_uid(structure_key(graph)) + int(run)))


def _has_edge(graph, a, b):
    return [a, b] in graph.get_list_edges(return_weights=False)


def _climb_confounders(graph, scorer):
    """The sequential algorithm of CGNN_confounders.py:219-383 against an abstract scorer (whose scores
    already include the per-edge complexity penalty, :237,275,303,342,362)."""
    visited = [graph.get_dict_nw()]
    best = scorer(graph)
    moved = True
    while moved:
        moved = False
        for link in graph.skeleton.get_list_edges_without_duplicate():
            _say(link)
            _say(graph.get_list_edges(return_weights=False))
            if _has_edge(graph, link[0], link[1]) or _has_edge(graph, link[1], link[0]):
                src, dst = (link[0], link[1]) if _has_edge(graph, link[0], link[1]) else (link[1], link[0])
                # ---- try the other orientation (:262-287)
                trial = deepcopy(graph)
                trial.reverse_edge(src, dst)
                shape = trial.get_dict_nw()
                if trial.is_cyclic() or shape in visited:
                    _say("No evaluation for edge " + str(src) + " -> " + str(dst))
                else:
                    _say("Reverse Edge " + str(src) + " -> " + str(dst) + " in evaluation")
                    visited.append(shape)
                    score = scorer(trial)
                    _say("Current score : " + str(score))
                    _say("Best score : " + str(best))
                    if score < best:
                        graph.reverse_edge(src, dst)
                        moved = True
                        best = score
                        _say("Edge " + str(src) + "->" + str(dst) + " got reversed !")
                        src, dst = dst, src
                # ---- try without the edge (:289-317)
                trial = deepcopy(graph)
                trial.remove_edge(src, dst)
                shape = trial.get_dict_nw()
                if shape in visited:
                    _say("Removing already evaluated for edge " + str(src) + " -> " + str(dst))
                else:
                    _say("Removing edge " + str(src) + " -> " + str(dst) + " in evaluation")
                    visited.append(shape)
                    score = scorer(trial)
                    _say("Current score : " + str(score))
                    _say("Best score : " + str(best))
                    if score < best:
                        graph.remove_edge(src, dst)
                        moved = True
                        best = score
                        _say("Edge " + str(src) + " -> " + str(dst) + " got removed, possible confounder !")
                    else:      # kept: its weight becomes the cost of removing it (:314-317)
                        _say("Edge " + str(src) + " -> " + str(dst) + " not removed. Score edge : " + str(score - best))
                        graph.set_weight(src, dst, score - best)
            else:
                # ---- a removed edge may come back, in either direction (:320-381)
                a, b = link[0], link[1]
                gains = []
                for u, v in ((a, b), (b, a)):
                    trial = deepcopy(graph)
                    trial.add(u, v)
                    shape = trial.get_dict_nw()
                    score = NO_SCORE
                    if trial.is_cyclic() or shape in visited:
                        _say("No addition possible for " + str(u) + " -> " + str(v))
                    else:
                        _say("Addition of edge " + str(u) + " -> " + str(v) + " in evaluation :")
                        visited.append(shape)
                        score = scorer(trial)
                        _say("score network add edge " + str(u) + " -> " + str(v) + " : " + str(score))
                    gains.append(score)
                _say("Best score : " + str(best))
                fwd, bwd = gains
                if fwd < best and fwd < bwd:
                    graph.add(a, b, best - fwd)
                    moved = True
                    _say("Edge " + str(a) + " -> " + str(b) + " is added with score : " + str(best - fwd) + " !")
                    best = fwd
                elif bwd < best and bwd < fwd:
                    graph.add(b, a, best - bwd)
                    moved = True
                    _say("Edge " + str(b) + " -> " + str(a) + " is added with score : " + str(best - bwd) + " !")
                    best = bwd
                else:
                    _say("Edge not added, possible confounder " + str(a) + " <-> " + str(b))
        scorer.end_of_sweep()
    return graph


def hill_climbing_confounders(graph, data, run_cgnn_function=run_CGNN_confounders_b200, **kwargs):
    """Greedy search over reverse / remove / add moves along the skeleton (CGNN_confounders.py:219-383).
    `graph` (a DirectedGraph with `.skeleton`) is improved in place and returned."""
    if run_cgnn_function is run_CGNN_confounders_b200:
        scorer = GraphScorer(data, confounders=True, **kwargs)
        final = run_search(_climb_confounders, graph, scorer)
        graph._graph = final._graph
        hill_climbing_confounders.last_batches = list(scorer.batches)
        hill_climbing_confounders.last_kernel_seconds = scorer.kernel_ms * 1e-3
        return graph
    return _climb_confounders(graph, _FunctionScorer(data, run_cgnn_function, penalty=True, **kwargs))


def exploratory_hill_climbing(graph, data, run_cgnn_function=run_CGNN_confounders_b200, **kwargs):
    raise NotImplementedError("unrunnable in the reference (CGNN_confounders.py:422-423); out of scope")


def tabu_search(graph, data, run_cgnn_function=run_CGNN_confounders_b200, **kwargs):
    raise ValueError("Not Yet Implemented")                      # CGNN_confounders.py:458


class CGNN_confounders(GraphModel):
    def __init__(self, backend="B200"):
        super(CGNN_confounders, self).__init__()
        if backend not in _BACKENDS:
            print("No backend known as {}".format(backend))
            raise ValueError
        self.backend = backend
        self.infer_graph = run_CGNN_confounders_b200

    def create_graph_from_data(self, data):
        print("The CGNN model is not able (yet?) to model the graph directly from raw data")
        raise ValueError

    def orient_directed_graph(self, data, dag, alg="HC", **kwargs):
        """Improve a DAG (with skeleton) under the hidden-variable model (CGNN_confounders.py:487-498)."""
        data = DataFrame(scale(data.values), columns=data.columns)
        algorithms = {"HC": hill_climbing_confounders, "tabu": tabu_search, "EHC": exploratory_hill_climbing}
        return algorithms[alg](dag, data, self.infer_graph, **kwargs)

    def orient_undirected_graph(self, data, umg, **kwargs):
        from .GNN import GNN
        dag = GNN(backend=self.backend).orient_graph_confounders(data, umg, **kwargs)
        return self.orient_directed_graph(data, dag, **kwargs)
