"""End-to-end decisions on the bundled example CSVs (BASELINE configs[0..2]) through the public
Python API.  The reference is unseeded, so the bar is decision-level: the pairwise directions must
equal Example_pairwise_targets.csv, and the graph paths must produce acyclic orientations of the
skeleton whose score does not get worse under hill climbing.  Reduced epochs keep the GPU test run
short; the full-length runs are recorded in profiles/ by tools/run_examples.py."""
import os

import numpy as np
import pandas as pd
import pytest

pytestmark = pytest.mark.gpu
EX = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "examples")


@pytest.fixture()
def settings():
    import cgnn
    from cgnn_b200 import runtime
    old = cgnn.SETTINGS.as_dict()
    cgnn.SETTINGS.verbose = False
    yield cgnn.SETTINGS
    for k, v in old.items():
        setattr(cgnn.SETTINGS, k, v)
    runtime.close_engines()


def test_pairwise_example_directions(settings, tmp_path):
    import cgnn
    from cgnn.utils import CCEPC_PairsFileReader as CC
    settings.h_layer_dim, settings.NB_RUNS = 30, 12
    settings.train_epochs, settings.test_epochs = 500, 200
    data = CC(os.path.join(EX, "Example_pairwise_pairs.csv"), scale=True)
    pred = cgnn.GNN(backend="TensorFlow").predict_dataset(data, printout=str(tmp_path / "p.csv"))
    target = pd.read_csv(os.path.join(EX, "Example_pairwise_targets.csv"))["Target"].values
    assert np.all(np.isfinite(pred)) and np.all(np.abs(pred) < 1)
    assert list(np.sign(pred)) == list(target), (pred, target)
    assert list(pd.read_csv(tmp_path / "p.csv").columns) == ["SampleID", "Predictions"]


def test_graph_example_orientation_and_hill_climbing(settings):
    import cgnn
    from cgnn_b200.CGNN import hill_climbing
    settings.NB_RUNS, settings.train_epochs, settings.test_epochs = 6, 150, 50
    data = pd.read_csv(os.path.join(EX, "Example_graph_numdata.csv"))
    umg = cgnn.UndirectedGraph(pd.read_csv(os.path.join(EX, "Example_graph_skeleton.csv")))
    dag = cgnn.GNN(backend="TensorFlow").orient_graph(data, umg)
    assert len(dag.get_list_edges()) == 30 and not dag.is_cyclic()
    skel = {frozenset(e) for e in umg.get_list_edges_without_duplicate()}
    assert {frozenset(e[:2]) for e in dag.get_list_edges()} == skel
    truth = {tuple(r) for r in pd.read_csv(os.path.join(EX, "Example_graph_target.csv")).values.tolist()}
    hits = sum((a, b) in truth for a, b, _ in dag.get_list_edges())
    assert hits >= 18, "pairwise orientation should beat a coin on the 30 edges (got %d)" % hits
    out = cgnn.CGNN(backend="TensorFlow").orient_directed_graph(data, dag)
    assert out is dag and not out.is_cyclic() and len(out.get_list_edges()) == 30
    assert max(hill_climbing.last_batches) > 1, "hill climbing must evaluate candidates in batches"


def test_confounder_example_runs(settings):
    import cgnn
    from cgnn_b200.utils.scaling import scale
    settings.NB_RUNS, settings.train_epochs, settings.test_epochs = 4, 100, 40
    data = pd.read_csv(os.path.join(EX, "Example_graph_confounders_numdata.csv"))
    data = pd.DataFrame(scale(data.values), columns=data.columns)
    skel = cgnn.UndirectedGraph(pd.read_csv(os.path.join(EX, "Example_graph_confounders_skeleton.csv")))
    dag = cgnn.GNN(backend="TensorFlow").orient_graph_confounders(data, skel)
    assert not dag.is_cyclic() and dag.skeleton is skel
    out = cgnn.CGNN_confounders(backend="TensorFlow").orient_directed_graph(data, dag)
    assert not out.is_cyclic() and 1 <= len(out.get_list_edges()) <= 24
    assert all(np.isfinite(w) for _, _, w in out.get_list_edges())


@pytest.mark.parametrize("conf", [False, True])
def test_batched_search_equals_one_by_one_on_the_gpu(settings, conf):
    """The batched, speculative search and a strictly sequential search (one candidate, one run per engine call,
    through the reference-style per-instance function) take the same decisions with the REAL engine: a
    candidate's score depends only on (seed, graph structure)."""
    import copy
    import cgnn
    from cgnn_b200.CGNN import hill_climbing, run_CGNN_b200
    from cgnn_b200.CGNN_confounders import hill_climbing_confounders, run_CGNN_confounders_b200
    rng = np.random.default_rng(3)
    n, names = 160, ["A", "B", "C", "D", "E"]
    a = rng.normal(size=n); b = np.tanh(a) + 0.3 * rng.normal(size=n); c = b ** 2 + 0.3 * rng.normal(size=n)
    d = rng.normal(size=n); e = c * 0.5 + d + 0.2 * rng.normal(size=n)
    df = pd.DataFrame(np.stack([a, b, c, d, e], 1), columns=names)
    df = (df - df.mean()) / df.std(ddof=0)
    edges = [("B", "A"), ("B", "C"), ("E", "C"), ("D", "E")]          # partly wrong orientations to leave room for moves
    skel = cgnn.UndirectedGraph(pd.DataFrame(edges + [("A", "D")], columns=["Cause", "Effect"])) if conf else None
    g = cgnn.DirectedGraph(skeleton=skel) if conf else cgnn.DirectedGraph()
    for k, (x, y) in enumerate(edges):
        g.add(x, y, 0.1 + 0.1 * k)
    kw = dict(nb_runs=3, h_layer_dim=8, train_epochs=40, test_epochs=10)
    climb, one = (hill_climbing_confounders, run_CGNN_confounders_b200) if conf else (hill_climbing, run_CGNN_b200)

    def per_instance(data, graph, idx, run, **k):                     # a distinct function object -> sequential driver
        return one(data, graph, idx, run, **k)

    g_bat = climb(copy.deepcopy(g), df, **kw)
    g_seq = climb(copy.deepcopy(g), df, per_instance, **kw)
    assert g_bat.get_dict_nw() == g_seq.get_dict_nw()
    for (x, y, w), (x2, y2, w2) in zip(g_bat.get_list_edges(), g_seq.get_list_edges()):
        assert (x, y) == (x2, y2) and abs(w - w2) <= 1e-6 * max(1.0, abs(w))
    assert not g_bat.is_cyclic()


def test_confounder_shape_full_step(engine):
    """The shapes of the bundled hidden-variable example: 500 rows, 19 variables, 24 shared noise columns."""
    from cgnn_b200.plan import build_plan, n_params
    from cgnn_b200.utils.scaling import scale
    from oracle import cgnn as oc
    import cgnn
    df = pd.read_csv(os.path.join(EX, "Example_graph_confounders_numdata.csv"))
    skel = cgnn.UndirectedGraph(pd.read_csv(os.path.join(EX, "Example_graph_confounders_skeleton.csv")))
    g = cgnn.DirectedGraph(skeleton=skel)
    for a, b in skel.get_list_edges_without_duplicate():
        g.add(a, b, 1.0)
    assert not g.is_cyclic()
    plan = build_plan(g, confounders=True)
    h = 20
    assert plan["n_noise"] == 19 + 24 and n_params(plan, h) == 2599
    data = scale(df[plan["list_nodes"]].values).astype(np.float32)
    rng = np.random.default_rng(0)
    theta = (0.05 * rng.normal(size=n_params(plan, h))).astype(np.float32)
    noise = rng.normal(size=(1, plan["n_noise"], 500)).astype(np.float32)
    got = engine.graph_step_debug(data, plan, noise, theta, h_layer_dim=h)
    ref = oc.step_debug(theta.astype(np.float64), data.astype(np.float64), noise[0].astype(np.float64),
                        oc.build_plan(g, confounders=True), h)
    rel = lambda x, y: float(np.max(np.abs(np.asarray(x, np.float64) - y)) / np.max(np.abs(y)))
    assert abs(got["loss"][0] - ref["loss"]) < 1e-5 * ref["loss"]
    assert rel(got["dldp"][0], ref["dloss_dpred"]) < 1e-5 and rel(got["grad"][0], ref["grad"]) < 1e-5
