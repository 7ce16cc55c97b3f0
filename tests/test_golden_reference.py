"""Pins the oracle restatement against fixtures produced by the reference's OWN model code
(Loss.py, GNN.py, CGNN.py, CGNN_confounders.py executed over oracle.tf1_shim by
oracle/make_golden.py): same inputs, same noise -> same losses at every sess.run and the same
.evaluate() mean, to fp64 round-off."""
import os

import numpy as np
import pandas as pd

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TIGHT = 1e-9


def graph_from_fixture(z, mod):
    skel = None
    if len(z["skeleton"]):
        skel = mod.UndirectedGraph(pd.DataFrame(z["skeleton"].tolist(), columns=["Cause", "Effect"]))
    g = mod.DirectedGraph(skeleton=skel) if skel is not None else mod.DirectedGraph()
    for (a, b), w in zip(z["edges"].tolist(), z["weights"].tolist()):
        g.add(a, b, w)
    return g


def test_mmd_loss_and_gradient():
    from oracle import mmd
    z = np.load(os.path.join(GOLD, "ref_shim_mmd.npz"))
    for k in range(3):
        t, p = z["t%d" % k], z["p%d" % k]
        loss, grad = mmd.mmd_loss_and_grad_autograd(t, p)
        assert abs(loss - float(z["loss%d" % k])) < 1e-12 * abs(loss)
        assert abs(mmd.mmd_loss_blocks(t, p) - float(z["loss%d" % k])) < TIGHT * abs(loss)
        np.testing.assert_allclose(mmd.mmd_grad_analytic(t, p), z["grad%d" % k], rtol=0, atol=TIGHT * np.abs(z["grad%d" % k]).max())


def test_pairwise_train_and_evaluate():
    from oracle import gnn
    z = np.load(os.path.join(GOLD, "ref_shim_pairwise.npz"))
    for tag in ("a", "b"):
        data, h, tr = z[tag + "_data"].astype(np.float64), int(z[tag + "_h"]), int(z[tag + "_train"])
        theta, losses, _ = gnn.train(z[tag + "_theta0"], data, z[tag + "_noise"][:tr], h)
        score, ev = gnn.evaluate(theta, data, z[tag + "_noise"][tr:], h)
        np.testing.assert_allclose(np.concatenate([losses, ev]), z[tag + "_losses"], rtol=TIGHT)
        assert abs(score - float(z[tag + "_score"])) < TIGHT * score


def test_pairwise_both_directions():
    """tf_run_instance (GNN.py:135-166): second model on the column-swapped data."""
    from oracle import gnn
    z = np.load(os.path.join(GOLD, "ref_shim_pairwise.npz"))
    data, h, tr = z["c_data"].astype(np.float64), int(z["c_h"]), int(z["c_train"])
    for d in range(2):
        dd = np.ascontiguousarray(data[:, ::-1]) if d else data
        s = gnn.run_instance(z["c_theta0"][d], dd, z["c_noise"][d, :tr], z["c_noise"][d, tr:], h)
        assert abs(s - z["c_scores"][d]) < TIGHT * s


def test_graph_and_confounder_models():
    import cgnn_b200.utils.Graph as my_graph
    from oracle import cgnn as oc
    for name, conf in (("graph", False), ("confounders", True)):
        z = np.load(os.path.join(GOLD, "ref_shim_%s.npz" % name))
        g = graph_from_fixture(z, my_graph)
        plan = oc.build_plan(g, confounders=conf)
        assert plan["list_nodes"] == z["list_nodes"].tolist()
        data, h, tr = z["data"].astype(np.float64), int(z["h"]), int(z["train"])
        assert oc.n_params(plan, h) == z["theta0"].size
        theta, losses, _ = oc.train(z["theta0"], data, z["noise"][:tr], plan, h)
        score, ev = oc.evaluate(theta, data, z["noise"][tr:], plan, h)
        np.testing.assert_allclose(np.concatenate([losses, ev]), z["losses"], rtol=TIGHT)
        assert abs(score - float(z["score"])) < TIGHT * score


def test_fourier_loss_and_fast_mmd_generator():
    """Fourier_MMD_Loss_tf (Loss.py:35-56) and GNN_tf(use_Fast_MMD=True) of the reference, with the random
    matrix its rp() drew: the oracle's restatement gives the same numbers."""
    import torch
    from oracle import mmd, gnn
    z = np.load(os.path.join(GOLD, "ref_shim_mmd.npz"))
    for k in range(2):
        v = float(mmd.fourier_mmd_literal(torch.tensor(z["ft%d" % k]), torch.tensor(z["fp%d" % k]), torch.tensor(z["fwz%d" % k]),
                                          int(z["fk%d" % k])))
        assert abs(v - float(z["floss%d" % k])) < 1e-12 * v
    z = np.load(os.path.join(GOLD, "ref_shim_pairwise.npz"))
    tr, k, h = int(z["f_train"]), int(z["f_k"]), int(z["f_h"])
    data = z["f_data"].astype(np.float64)
    theta, losses, _ = gnn.train(z["f_theta0"], data, z["f_noise"][:tr], h, wz=z["f_wz"][:tr], k=k)
    score, ev = gnn.evaluate(theta, data, z["f_noise"][tr:], h, wz=z["f_wz"][tr:], k=k)
    np.testing.assert_allclose(np.concatenate([losses, ev]), z["f_losses"], rtol=TIGHT)
    assert abs(score - float(z["f_score"])) < TIGHT * score


def run_instance_case(z, tag, graph_mod):
    """(plan, data, noise in the engine's column ids, theta0, h, tr, te) of one case of ref_shim_run_instance.npz.
    The plan comes from cgnn_b200.plan.build_plan with the DataFrame's (shuffled) column order."""
    from cgnn_b200.plan import build_plan
    conf = tag == "conf_"
    sub = {k: z[tag + k] for k in ("edges", "weights", "skeleton")}
    g = graph_from_fixture(sub, graph_mod)
    plan = build_plan(g, columns=z[tag + "columns"].tolist(), confounders=conf)
    own, shared = z[tag + "own_noise"], z[tag + "shared_noise"]
    steps, n_var, n = own.shape
    noise = np.zeros((steps, plan["n_noise"], n))
    for pos, nd in enumerate(plan["nodes"]):
        noise[:, nd["noise"][0]] = own[:, pos]
    for e in range(shared.shape[1]):
        noise[:, n_var + e] = shared[:, e]
    return plan, z[tag + "data"], noise, z[tag + "theta0"], int(z[tag + "h"]), int(z[tag + "train"]), int(z[tag + "test"])


def test_run_instance_entry_points_with_shuffled_columns():
    """run_CGNN_tf / run_CGNN_confounders_tf of the reference on a DataFrame whose columns are not in node order:
    the host layer's own column mapping (cgnn_b200/plan.py) must lead the oracle to the same numbers."""
    import cgnn_b200.utils.Graph as my_graph
    from oracle import cgnn as oc
    z = np.load(os.path.join(GOLD, "ref_shim_run_instance.npz"))
    for tag in ("plain_", "conf_"):
        plan, data, noise, theta0, h, tr, te = run_instance_case(z, tag, my_graph)
        assert plan["columns"] != plan["list_nodes"]
        theta, losses, _ = oc.train(theta0, data.astype(np.float64), noise[:tr], plan, h)
        score, ev = oc.evaluate(theta, data.astype(np.float64), noise[tr:], plan, h)
        np.testing.assert_allclose(np.concatenate([losses, ev]), z[tag + "losses"], rtol=TIGHT)
        assert abs(score - float(z[tag + "score"])) < TIGHT * score


def test_graph_generator_with_the_random_feature_loss():
    """CGNN_tf(use_Fast_MMD=True) of the reference (CGNN.py:92-93): losses of every train and evaluate sess.run
    and the evaluate() mean, replayed by the oracle with the recorded noise and rp() matrices."""
    import cgnn_b200.utils.Graph as my_graph
    from oracle import cgnn as oc
    from oracle.adam import TFAdam
    z = np.load(os.path.join(GOLD, "ref_shim_graph_fourier.npz"))
    g = graph_from_fixture(z, my_graph)
    plan = oc.build_plan(g)
    assert plan["list_nodes"] == z["list_nodes"].tolist()
    data, h, tr, te, k = z["data"].astype(np.float64), int(z["h"]), int(z["train"]), int(z["test"]), int(z["k"])
    theta = z["theta0"].astype(np.float64)
    opt = TFAdam(theta.size, 0.01)
    losses = []
    for s in range(tr + te):
        r = oc.step_debug(theta, data, z["noise"][s], plan, h, wz=z["wz"][s], k=k)
        losses.append(r["loss"])
        if s < tr:
            theta = opt.step(theta, r["grad"])
    np.testing.assert_allclose(losses, z["losses"], rtol=TIGHT)
    assert abs(np.mean(losses[tr:]) - float(z["score"])) < TIGHT * float(z["score"])


def test_rp_literal_shape_and_scales():
    from oracle.mmd import rp_literal
    from oracle import GAMMAS
    wz = rp_literal(500, 3, np.random.RandomState(0))
    assert wz.shape == (4, 3500)
    for g, gam in enumerate(GAMMAS):
        blk = wz[:3, 500 * g:500 * (g + 1)]
        assert abs(blk.std() / (2 * gam) - 1) < 0.1
    assert 0 <= wz[3].min() and wz[3].max() <= 2 * np.pi
