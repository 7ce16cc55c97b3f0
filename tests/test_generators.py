"""Synthetic data helpers (host side): deterministic, well-formed, readable by the pairs reader."""
import numpy as np


def test_random_graph_generator_roundtrip(tmp_path):
    from cgnn_b200.generators import RandomGraphGenerator
    from cgnn_b200.utils import CCEPC_PairsFileReader
    g1 = RandomGraphGenerator(num_nodes=15, max_joint_causes=3, number_points=120, seed=3)
    graph, data, cat, cat_idx = g1.generate()
    assert data.shape == (120, 15) and not graph.is_cyclic() and cat.shape == data.shape and len(cat_idx) == 3
    assert np.allclose(data.iloc[:, 5:].values.mean(0), 0, atol=1e-9) and np.allclose(data.iloc[:, 5:].values.std(0), 1, atol=1e-9)
    assert set(graph.get_list_nodes()) <= set(data.columns)
    g2 = RandomGraphGenerator(num_nodes=15, max_joint_causes=3, number_points=120, seed=3)
    assert g2.generate()[1].equals(data)                       # seeded
    g1.save_data(str(tmp_path / "syn"))
    assert (tmp_path / "syn_numdata.csv").exists() and (tmp_path / "syn_target.csv").exists()
    pairs, targets = g1.generate_pairs(7)
    pairs.to_csv(tmp_path / "pairs.csv", index=False)
    back = CCEPC_PairsFileReader(str(tmp_path / "pairs.csv"), scale=False)
    assert len(back) == 7 and len(back["A"][0]) == 120 and set(targets["Target"]) <= {1.0, -1.0}


def test_synth_pairs_and_dag_are_deterministic():
    from cgnn_b200 import synth
    a, la = synth.make_pairs(10, 3, 200)
    b, lb = synth.make_pairs(10, 3, 200)
    assert np.array_equal(a, b) and np.array_equal(la, lb) and a.dtype == np.float32
    assert np.allclose(a.mean(1), 0, atol=1e-5) and np.allclose(a.std(1), 1, atol=1e-4)
    X, edges = synth.make_dag_dataset(12, 300, 3, seed=1)
    assert X.shape == (300, 12) and all(p < c for p, c in edges)
