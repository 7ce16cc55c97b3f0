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
