"""cgnn_b200.utils.Graph against the reference's utils/Graph.py (which runs unmodified here) and
against frozen fixtures (tests/golden/graph_ops.json) for machines without /root/reference."""
import json
import os

import numpy as np
import pandas as pd
import pytest

import refload

HERE = os.path.dirname(os.path.abspath(__file__))
EX = os.path.join(os.path.dirname(HERE), "examples")
GOLD = os.path.join(HERE, "golden", "graph_ops.json")


def scenario(mod, seed):
    """Runs a scripted sequence of operations on module `mod`'s classes; returns a JSON-able trace."""
    rng = np.random.default_rng(seed)
    n = int(rng.integers(4, 9))
    names = ["V%d" % k for k in rng.permutation(n)]
    g = mod.DirectedGraph()
    trace = []
    for _ in range(int(rng.integers(n, 3 * n))):
        a, b = rng.choice(n, 2, replace=False)
        w = float(np.round(rng.uniform(0, 1), 3)) if rng.uniform() > 0.2 else 0.5     # ties on purpose
        g.add(names[a], names[b], w)
    snap = lambda: dict(nodes=g.get_list_nodes(), edges=g.get_list_edges(), desc=g.get_list_edges(descending=True),
                        plain=g.get_list_edges(order_by_weight=False, return_weights=False), nw=g.get_dict_nw(),
                        parents={v: g.get_parents(v) for v in g.get_list_nodes()}, cyclic=g.is_cyclic(),
                        ncycles=len(g.cycles()), first_cycle=(g.cycles() or [None])[0])
    trace.append(snap())
    import copy
    h = copy.deepcopy(g)
    h.remove_cycle_without_deletion()
    trace.append(dict(rcwd=h.get_list_edges(order_by_weight=False), cyclic=h.is_cyclic(), nodes=h.get_list_nodes()))
    import contextlib, io
    with contextlib.redirect_stdout(io.StringIO()):
        g.remove_cycles()
    trace.append(snap())
    edges = g.get_list_edges(order_by_weight=False, return_weights=False)
    for k in range(min(3, len(edges))):
        a, b = edges[(k * 5) % len(edges)]
        if [a, b] in g.get_list_edges(order_by_weight=False, return_weights=False):
            g.reverse_edge(a, b)
            trace.append(snap())
    if g.get_list_edges(order_by_weight=False):
        a, b = g.get_list_edges(order_by_weight=False, return_weights=False)[0]
        g.set_weight(a, b, 0.123)
        g.remove_edge(*g.get_list_edges(order_by_weight=False, return_weights=False)[-1])
        trace.append(snap())
    m, nodes = g.get_adjacency_matrix()
    trace.append(dict(adj=np.asarray(m).tolist(), adj_nodes=nodes))
    return json.loads(json.dumps(trace))


def skeleton_trace(mod):
    out = {}
    for f in ("Example_graph_skeleton.csv", "Example_graph_confounders_skeleton.csv"):
        u = mod.UndirectedGraph(pd.read_csv(os.path.join(EX, f)))
        out[f] = dict(nodes=u.get_list_nodes(), edges=u.get_list_edges_without_duplicate(),
                      neigh={v: u.get_neighbors(v) for v in u.get_list_nodes()},
                      all_edges=u.get_list_edges(order_by_weight=False, return_weights=False))
        a, b = u.get_list_edges_without_duplicate()[3]
        u.remove_edge(a, b)
        out[f + ":removed"] = dict(nodes=u.get_list_nodes(), edges=u.get_list_edges_without_duplicate())
    d = mod.DirectedGraph(pd.read_csv(os.path.join(EX, "Example_graph_target.csv")))
    out["target"] = dict(nodes=d.get_list_nodes(), cyclic=d.is_cyclic(), nw=d.get_dict_nw(),
                         parents={v: d.get_parents(v) for v in d.get_list_nodes()})
    adj = pd.DataFrame([[0, 0.5, 0], [0, 0, 0.7], [0.0005, 0, 0]], columns=["a", "b", "c"])
    out["adjacency"] = mod.DirectedGraph(adj, adjacency_matrix=True).get_list_edges()
    return json.loads(json.dumps(out))


def test_graph_ops_match_reference_and_golden():
    import cgnn_b200.utils.Graph as mine
    got = dict(scenarios=[scenario(mine, s) for s in range(40)], skeletons=skeleton_trace(mine))
    if refload.available():
        ref = refload.load()["Graph"]
        want = dict(scenarios=[scenario(ref, s) for s in range(40)], skeletons=skeleton_trace(ref))
        assert got == want
        if os.environ.get("CGNN_WRITE_GOLDEN"):
            json.dump(want, open(GOLD, "w"))
    assert os.path.exists(GOLD), "golden fixture missing: run tests/make_host_golden.sh in the build container"
    assert got == json.load(open(GOLD))


def test_empty_graph_is_handled():
    from cgnn_b200.utils.Graph import DirectedGraph
    g = DirectedGraph()
    assert g.get_list_edges() == [] and g.get_list_nodes() == [] and not g.is_cyclic()


def test_bundled_target_dag_facts():
    """Facts the survey measured with the reference's own Graph.py."""
    from copy import deepcopy
    from cgnn_b200.utils.Graph import DirectedGraph
    from cgnn_b200.plan import build_plan, n_params
    d = DirectedGraph(pd.read_csv(os.path.join(EX, "Example_graph_target.csv")))
    assert len(d.get_list_edges()) == 30 and not d.is_cyclic()
    acyclic = 0
    for a, b, _ in d.get_list_edges():
        t = deepcopy(d)
        t.reverse_edge(a, b)
        acyclic += not t.is_cyclic()
    assert acyclic == 27
    plan = build_plan(d)
    assert sum(len(nd["parents"]) + 1 for nd in plan["nodes"]) == 52 and n_params(plan, 20) == 1942
