"""Host logic of the pairwise path without a GPU: CSV reader, standardisation, batching, score
formula, orientation of a skeleton - against the reference's own Python where it runs here."""
import os

import numpy as np
import pandas as pd
import pytest

import refload
from fake_engine import FakeEngine, pair_run_score

HERE = os.path.dirname(os.path.abspath(__file__))
EX = os.path.join(os.path.dirname(HERE), "examples")


@pytest.fixture()
def fake(monkeypatch):
    from cgnn_b200 import runtime, SETTINGS
    eng = FakeEngine()
    monkeypatch.setattr(runtime, "get_engines", lambda: [eng])
    old = SETTINGS.verbose
    SETTINGS.verbose = False
    yield eng
    SETTINGS.verbose = old


def test_pairs_reader_matches_reference():
    from cgnn_b200.utils import CCEPC_PairsFileReader
    f = os.path.join(EX, "Example_pairwise_pairs.csv")
    for sc in (True, False):
        mine = CCEPC_PairsFileReader(f, scale=sc)
        assert list(mine.columns) == ["SampleID", "A", "B"] and len(mine) == 5 and len(mine["A"][0]) == 1500
        if sc:
            assert abs(mine["A"][2].mean()) < 1e-12 and abs(mine["B"][4].std() - 1) < 1e-12
        if refload.available():
            ref = refload.load()["Formats"].CCEPC_PairsFileReader(f, scale=sc)
            assert list(ref["SampleID"]) == list(mine["SampleID"])
            for col in ("A", "B"):
                for x, y in zip(ref[col], mine[col]):
                    np.testing.assert_allclose(x, y, rtol=0, atol=1e-12)


def test_scale_matches_sklearn():
    from sklearn.preprocessing import scale as sk
    from cgnn_b200.utils.scaling import scale
    rng = np.random.default_rng(0)
    x = rng.normal(size=(200, 4)) * [1, 10, 0.01, 3] + [0, 5, -2, 100]
    x[:, 2] = 7.0                                   # constant column: centred only
    np.testing.assert_allclose(scale(x), sk(x), atol=1e-12)
    np.testing.assert_allclose(scale(x[:, 1]), sk(x[:, 1]), atol=1e-12)


def test_predict_dataset_batches_and_scores(fake, tmp_path):
    import cgnn_b200 as cgnn
    from cgnn_b200.utils import CCEPC_PairsFileReader
    df = CCEPC_PairsFileReader(os.path.join(EX, "Example_pairwise_pairs.csv"))
    out = tmp_path / "printout.csv"
    pred = cgnn.GNN(backend="TensorFlow").predict_dataset(df, printout=str(out), nb_runs=3)
    assert len(fake.calls) == 1 and fake.calls[0][:3] == ("pairwise", 5, 3)      # ONE submission for the dataset
    # the score formula of GNN.py:196-207 on the fake per-run scores
    from cgnn_b200.utils.scaling import scale
    for k in range(5):
        a = scale(np.asarray(df["A"][k]).reshape(-1, 1)).astype(np.float32)[:, 0]
        b = scale(np.asarray(df["B"][k]).reshape(-1, 1)).astype(np.float32)[:, 0]
        w = np.arange(1500) % 7
        ab = np.mean([np.float32(pair_run_score((k * 3 + r) * 2 + 0, float(np.dot(a, w)))) for r in range(3)])
        ba = np.mean([np.float32(pair_run_score((k * 3 + r) * 2 + 1, float(np.dot(b, w)))) for r in range(3)])
        assert abs(pred[k] - (ba - ab) / (ba + ab)) < 1e-6
    got = pd.read_csv(out)
    assert list(got.columns) == ["SampleID", "Predictions"] and list(got["SampleID"]) == list(df["SampleID"])
    np.testing.assert_allclose(got["Predictions"], pred)
    # predict_proba of a single pair with its idx reproduces the batched value (uid = idx-keyed)
    one = cgnn.GNN().predict_proba(df["A"][3], df["B"][3], 3, nb_runs=3)
    a3 = scale(np.asarray(df["A"][3]).reshape(-1, 1))
    b3 = scale(np.asarray(df["B"][3]).reshape(-1, 1))
    assert abs(cgnn.GNN().predict_proba(a3, b3, 3, nb_runs=3) - pred[3]) < 1e-12 and np.isfinite(one)


def test_predictions_equal_the_reference_code_on_scripted_run_scores(fake, capsys):
    """The reference's own GNN.predict_proba / predict_dataset (GNN.py:180-207, PairwiseModel.py:32-55), with
    tf_run_instance replaced by the fake engine's per-run score function, against the host layer driven by the
    fake engine: same standardisation, same averaging over runs, same (BA-AB)/(BA+AB)."""
    if not refload.available():
        pytest.skip("needs the reference tree (build container only)")
    import cgnn_b200 as cgnn
    from cgnn_b200.utils import CCEPC_PairsFileReader
    ref = refload.load()
    nb_runs = 4

    def scripted_run_instance(m, idx, run, **kw):          # signature of tf_run_instance (GNN.py:135)
        w = np.arange(m.shape[0]) % 7
        uid = (idx * nb_runs + run) * 2
        return [np.float32(pair_run_score(uid, float(np.dot(m[:, 0], w)))), np.float32(pair_run_score(uid + 1, float(np.dot(m[:, 1], w))))]

    ref["GNN"].tf_run_instance = scripted_run_instance
    ref["Settings"].SETTINGS.NB_RUNS, ref["Settings"].SETTINGS.NB_JOBS = nb_runs, 1     # predict_dataset takes no kwargs there
    f = os.path.join(EX, "Example_pairwise_pairs.csv")
    rdf = ref["Formats"].CCEPC_PairsFileReader(f, scale=True)
    want = ref["GNN"].GNN(backend="TensorFlow").predict_dataset(rdf)
    capsys.readouterr()
    got = cgnn.GNN(backend="TensorFlow").predict_dataset(CCEPC_PairsFileReader(f, scale=True), nb_runs=nb_runs)
    np.testing.assert_allclose(got, np.asarray(want, dtype=np.float64), rtol=0, atol=2e-7)
    one = ref["GNN"].GNN(backend="TensorFlow").predict_proba(rdf["A"][2], rdf["B"][2], 2, nb_runs=nb_runs, nb_jobs=1)
    capsys.readouterr()
    assert abs(cgnn.GNN().predict_proba(rdf["A"][2], rdf["B"][2], 2, nb_runs=nb_runs) - one) < 2e-7


def test_unknown_backend_raises():
    import cgnn_b200 as cgnn
    with pytest.raises(KeyError):
        cgnn.GNN(backend="Caffe")
    with pytest.raises(ValueError):
        cgnn.CGNN(backend="Caffe")
    with pytest.raises(ValueError):
        cgnn.CGNN().create_graph_from_data(None)
    with pytest.raises(ValueError):
        cgnn.CGNN().predict(None, graph=3)


def test_long_pairs_are_subsampled_per_run(fake):
    """GNN.py:150-153: more than max_nb_points rows -> one random subset per run."""
    import cgnn_b200 as cgnn
    rng = np.random.default_rng(0)
    a, b = rng.normal(size=4000), rng.normal(size=4000)
    s = cgnn.GNN().predict_proba(a, b, 0, nb_runs=4)
    assert np.isfinite(s)
    kind, n_slots, runs_per_slot, _ = fake.calls[0]
    assert (n_slots, runs_per_slot) == (4, 1)


class _Scripted(object):
    """A pairwise model whose score is a fixed function of the two columns (works for both code bases)."""
    def predict_proba(self, a, b, idx=0, **kw):
        return float(np.tanh(np.mean(np.asarray(a) ** 3) - np.mean(np.asarray(b) ** 3)))


@pytest.mark.parametrize("conf", [False, True])
def test_orient_graph_matches_reference_flow(conf, capsys):
    """orient_graph / orient_graph_confounders (PairwiseModel.py:57-125) with a scripted predict_proba:
    same edges, weights and cycle handling as the reference's own base class."""
    import cgnn_b200.utils.Graph as my_graph
    from cgnn_b200.PairwiseModel import Pairwise_Model
    data = pd.read_csv(os.path.join(EX, "Example_graph_confounders_numdata.csv" if conf else "Example_graph_numdata.csv"))
    skel = pd.read_csv(os.path.join(EX, "Example_graph_confounders_skeleton.csv" if conf else "Example_graph_skeleton.csv"))

    class Mine(_Scripted, Pairwise_Model):
        pass

    m = Mine()
    umg = my_graph.UndirectedGraph(skel)
    g = m.orient_graph_confounders(data, umg) if conf else m.orient_graph(data, umg)
    assert not g.is_cyclic()
    if refload.available():
        ref = refload.load()

        class Ref(_Scripted, ref["PairwiseModel"].Pairwise_Model):
            pass

        r = Ref()
        rumg = ref["Graph"].UndirectedGraph(skel)
        rg = r.orient_graph_confounders(data, rumg) if conf else r.orient_graph(data, rumg)
        capsys.readouterr()
        assert rg.get_dict_nw() == g.get_dict_nw()
        for (a, b, w), (a2, b2, w2) in zip(rg.get_list_edges(), g.get_list_edges()):
            assert (a, b) == (a2, b2) and abs(w - w2) < 1e-12


def test_run_instance_seam_equals_the_batched_path_for_long_pairs(fake):
    """A pair longer than SETTINGS.max_nb_points is cut to a fresh random row subset per run (GNN.py:150-153).  The
    per-instance seam (drop-in for tf_run_instance) must draw the subset of ITS (pair, run) - not the same rows for every
    run - and so reproduce the batched scores."""
    from cgnn_b200 import SETTINGS
    from cgnn_b200.GNN import b200_run_instance, score_pairs
    rng = np.random.default_rng(3)
    m = rng.normal(size=(1700, 2)).astype(np.float32)
    assert m.shape[0] > SETTINGS.max_nb_points
    batched = score_pairs([m], 3, nb_runs=4)
    single = np.array([b200_run_instance(m, 3, run, nb_runs=4) for run in range(4)], np.float32)
    assert np.array_equal(batched[0], single)
    assert len({tuple(r) for r in single.tolist()}) == 4           # four different row subsets
