"""Host-side search logic (no GPU): the batched, speculative hill climbing must take exactly the
decisions of the reference's sequential code.  Three executions are compared on the same
deterministic fake scorer: (1) the REFERENCE's own hill_climbing / hill_climbing_confounders
(loaded read-only from /root/reference when present), (2) this repo's sequential path (custom
scorer function) and (3) this repo's batched path (memoised re-execution over a fake engine)."""
import json
import os

import numpy as np
import pandas as pd
import pytest

import refload
from fake_engine import FakeEngine, graph_run_score

HERE = os.path.dirname(os.path.abspath(__file__))
EX = os.path.join(os.path.dirname(HERE), "examples")
GOLD = os.path.join(HERE, "golden", "host_search.json")


def _key(graph):
    return tuple(sorted((str(a), str(b)) for a, b in graph.get_list_edges(order_by_weight=False, return_weights=False)))


def fake_fn(data, graph, idx, run, **kw):
    return graph_run_score(_key(graph), run)


def _pairwise_weight(a, b):
    from fake_engine import _h
    return 2.0 * _h("w", a, b) - 1.0


def build_start(mod_graph, conf):
    """Pairwise-oriented start graph from the bundled skeleton, with deterministic fake pairwise weights."""
    skel_file = "Example_graph_confounders_skeleton.csv" if conf else "Example_graph_skeleton.csv"
    umg = mod_graph.UndirectedGraph(pd.read_csv(os.path.join(EX, skel_file)))
    g = mod_graph.DirectedGraph(skeleton=umg) if conf else mod_graph.DirectedGraph()
    for a, b in umg.get_list_edges_without_duplicate():
        w = _pairwise_weight(a, b)
        if w > 0:
            g.add(a, b, w)
        else:
            g.add(b, a, abs(w))
    if conf:
        g.remove_cycles(verbose=False) if "verbose" in g.remove_cycles.__code__.co_varnames else g.remove_cycles()
    else:
        g.remove_cycle_without_deletion()
    return g


def _data(conf):
    f = "Example_graph_confounders_numdata.csv" if conf else "Example_graph_numdata.csv"
    return pd.read_csv(os.path.join(EX, f))


def _edges(g):
    return [[str(a), str(b)] for a, b in g.get_list_edges(order_by_weight=False, return_weights=False)]


@pytest.fixture()
def quiet():
    from cgnn_b200 import SETTINGS
    old = SETTINGS.verbose
    SETTINGS.verbose = False
    yield
    SETTINGS.verbose = old


@pytest.mark.parametrize("conf", [False, True])
def test_batched_search_equals_sequential_and_reference(conf, quiet, monkeypatch, capsys):
    import cgnn_b200.utils.Graph as my_graph
    from cgnn_b200 import runtime
    from cgnn_b200.CGNN import hill_climbing
    from cgnn_b200.CGNN_confounders import hill_climbing_confounders
    climb = hill_climbing_confounders if conf else hill_climbing
    data = _data(conf)
    nb_runs = 4

    # (2) sequential path of this repo, custom per-instance function
    g_seq = climb(build_start(my_graph, conf), data, fake_fn, nb_runs=nb_runs)
    # (3) batched path over a fake engine
    eng = FakeEngine()
    monkeypatch.setattr(runtime, "get_engines", lambda: [eng])
    g_bat = climb(build_start(my_graph, conf), data, nb_runs=nb_runs)
    assert _edges(g_bat) == _edges(g_seq)
    assert g_bat.get_dict_nw() == g_seq.get_dict_nw()
    w_seq = {(a, b): w for a, b, w in g_seq.get_list_edges(order_by_weight=False)}
    w_bat = {(a, b): w for a, b, w in g_bat.get_list_edges(order_by_weight=False)}
    assert w_seq.keys() == w_bat.keys() and all(abs(w_seq[k] - w_bat[k]) < 1e-6 for k in w_seq)
    n_batches = len(climb.last_batches)
    n_graphs = sum(climb.last_batches)
    assert n_batches < n_graphs, "candidates must be evaluated in batches, not one by one"

    # (1) the reference's own search code on the same scorer
    if refload.available():
        ref = refload.load()
        ref_climb = ref["CGNN_confounders"].hill_climbing_confounders if conf else ref["CGNN"].hill_climbing
        g_ref = ref_climb(build_start(ref["Graph"], conf), data, fake_fn, nb_runs=nb_runs, nb_jobs=1)
        capsys.readouterr()
        assert _edges(g_ref) == _edges(g_seq)
        assert g_ref.get_dict_nw() == g_seq.get_dict_nw()
        w_ref = {(a, b): w for a, b, w in g_ref.get_list_edges(order_by_weight=False)}
        assert all(abs(w_ref[k] - w_seq[k]) < 1e-9 for k in w_ref)

    # frozen result (checked everywhere, also where /root/reference is absent)
    gold = json.load(open(GOLD)) if os.path.exists(GOLD) else {}
    name = "confounders" if conf else "plain"
    if os.environ.get("CGNN_WRITE_GOLDEN"):
        gold[name] = _edges(g_seq)
        json.dump(gold, open(GOLD, "w"), indent=0)
    assert gold.get(name) == _edges(g_seq)


def test_search_accepts_some_moves(quiet):
    """The fake scorer must actually exercise acceptance (otherwise the equality above is hollow)."""
    import cgnn_b200.utils.Graph as my_graph
    from cgnn_b200.CGNN import hill_climbing
    start = build_start(my_graph, False)
    before = _edges(start)
    after = _edges(hill_climbing(start, _data(False), fake_fn, nb_runs=2))
    assert before != after


def test_confounder_search_with_a_missing_edge_uses_the_add_branch(quiet, monkeypatch):
    """Regression: while speculating, the 'not evaluated' sentinel 9999 (CGNN_confounders.py:330,349) must
    not win the add decision - a start graph with a removed edge exercises that branch."""
    import copy
    import cgnn_b200.utils.Graph as my_graph
    from cgnn_b200 import runtime
    from cgnn_b200.CGNN_confounders import hill_climbing_confounders
    monkeypatch.setattr(runtime, "get_engines", lambda: [FakeEngine()])
    data = _data(True)
    start = build_start(my_graph, True)
    for a, b, _ in start.get_list_edges()[:4]:
        g1 = copy.deepcopy(start)
        g1.remove_edge(a, b)
        g2 = copy.deepcopy(g1)
        bat = hill_climbing_confounders(g1, data, nb_runs=2)
        seq = hill_climbing_confounders(g2, data, fake_fn, nb_runs=2)
        assert not bat.is_cyclic() and _edges(bat) == _edges(seq)
        if refload.available():
            import contextlib, io
            ref = refload.load()
            g3 = build_start(ref["Graph"], True)
            g3.remove_edge(a, b)
            with contextlib.redirect_stdout(io.StringIO()):
                r = ref["CGNN_confounders"].hill_climbing_confounders(g3, data, fake_fn, nb_runs=2, nb_jobs=1)
            assert _edges(r) == _edges(seq)


def test_canonical_plan_ignores_insertion_history():
    """The search caches scores by edge set; the plan (hence the flat parameter layout and which initial weight each
    connection draws) must then be a function of the edge set alone - not of the order in which edges were added or
    reversed."""
    import pandas as pd
    from cgnn_b200.plan import build_plan
    from cgnn_b200.utils.Graph import DirectedGraph, UndirectedGraph
    edges = [("V3", "V1"), ("V0", "V1"), ("V1", "V2"), ("V0", "V4"), ("V4", "V2")]
    cols = ["V%d" % k for k in range(5)]
    skel = UndirectedGraph(pd.DataFrame(edges, columns=["Cause", "Effect"]))
    skel2 = UndirectedGraph(pd.DataFrame([(b, a) for a, b in reversed(edges)], columns=["Cause", "Effect"]))
    plans = []
    for order, sk in ((edges, skel), (list(reversed(edges)), skel2)):
        g = DirectedGraph(skeleton=sk)
        for a, b in order:
            g.add(a, b, 1.0)
        g.reverse_edge("V0", "V4"); g.reverse_edge("V4", "V0")          # same edge set again, different insertion history
        plans.append((build_plan(g, columns=cols, canonical=True), build_plan(g, columns=cols, confounders=True, canonical=True)))
    for a, b in zip(plans[0], plans[1]):
        assert a["nodes"] == b["nodes"] and a["n_noise"] == b["n_noise"]
    # the literal (reference-order) plans do differ - that is what canonical=True removes
    g1, g2 = DirectedGraph(), DirectedGraph()
    for a, b in edges:
        g1.add(a, b, 1.0)
    for a, b in reversed(edges):
        g2.add(a, b, 1.0)
    assert build_plan(g1, columns=cols)["nodes"] != build_plan(g2, columns=cols)["nodes"]


def test_verbose_search_prints_every_decision_once(monkeypatch, capsys):
    """The batched search re-executes the algorithm once per batch; with SETTINGS.verbose the reference's progress lines
    (CGNN.py:220,233-248) must come out once - from the final, complete execution - not once per re-execution."""
    from fake_engine import FakeEngine
    from cgnn_b200 import runtime, SETTINGS
    from cgnn_b200.CGNN import hill_climbing
    import cgnn_b200.utils.Graph as my_graph
    monkeypatch.setattr(runtime, "get_engines", lambda: [FakeEngine()])
    monkeypatch.setattr(SETTINGS, "verbose", True)
    hill_climbing(build_start(my_graph, False), _data(False), nb_runs=2)
    out = capsys.readouterr().out.splitlines()
    assert len(hill_climbing.last_batches) > 2
    assert sum(1 for l in out if l.startswith("Graph score")) == 1
    evaluated = [l for l in out if l.endswith("in evaluation :")]
    assert len(evaluated) == len(set(evaluated)) or len(evaluated) >= 1       # an edge may be re-tried in a later sweep, never re-printed per batch
    assert len(evaluated) <= sum(hill_climbing.last_batches)
