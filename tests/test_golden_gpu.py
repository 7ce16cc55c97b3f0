"""The CUDA kernels, through the C ABI, against the fixtures the reference's own model code
produced (tests/golden/ref_shim_*.npz, see oracle/make_golden.py): identical data, initial weights
and noise -> MMD values of every train / evaluate sess.run within 1e-5 relative, final scores too."""
import os

import numpy as np
import pytest

from test_golden_reference import GOLD, graph_from_fixture
from util import relerr

pytestmark = pytest.mark.gpu
TOL = 1e-5


@pytest.mark.parametrize("pol", [0, 1, 2, 3])
def test_pairwise_fixtures(engine, pol):
    z = np.load(os.path.join(GOLD, "ref_shim_pairwise.npz"))
    for tag in ("a", "b"):
        data, h, tr, te = z[tag + "_data"], int(z[tag + "_h"]), int(z[tag + "_train"]), int(z[tag + "_test"])
        n = data.shape[0]
        init = np.repeat(z[tag + "_theta0"][None], 2, 0).astype(np.float32)
        noise = np.repeat(z[tag + "_noise"][None], 2, 0).astype(np.float32)
        r = engine.run_pairwise(data[None], n_runs=1, noise=noise, init_params=init, want_trace=True, h_layer_dim=h,
                                train_epochs=tr, test_epochs=te, trace_every=1, exp_policy=pol)
        assert relerr(r["trace"][0], z[tag + "_losses"]) < 2 * TOL      # float32 inputs of a fp64 fixture
        assert abs(r["scores"][0, 0, 0] - float(z[tag + "_score"])) < 2 * TOL * float(z[tag + "_score"])
    # both directions, as tf_run_instance pairs them
    data, h, tr, te = z["c_data"], int(z["c_h"]), int(z["c_train"]), int(z["c_test"])
    r = engine.run_pairwise(data[None], n_runs=1, noise=z["c_noise"].astype(np.float32), init_params=z["c_theta0"].astype(np.float32),
                            h_layer_dim=h, train_epochs=tr, test_epochs=te, exp_policy=pol)
    assert relerr(r["scores"][0, 0], z["c_scores"]) < 2 * TOL


@pytest.mark.parametrize("pol", [0, 1, 3])
@pytest.mark.parametrize("name,conf", [("graph", False), ("confounders", True)])
def test_graph_fixtures(engine, name, conf, pol):
    import cgnn_b200.utils.Graph as my_graph
    from cgnn_b200.plan import build_plan
    z = np.load(os.path.join(GOLD, "ref_shim_%s.npz" % name))
    plan = build_plan(graph_from_fixture(z, my_graph), confounders=conf)
    h, tr, te = int(z["h"]), int(z["train"]), int(z["test"])
    r = engine.run_graph(z["data"], [plan], n_runs=1, noise=z["noise"][None].astype(np.float32),
                         init_params=z["theta0"][None].astype(np.float32), want_trace=True, h_layer_dim=h,
                         train_epochs=tr, test_epochs=te, trace_every=1, exp_policy=pol)
    assert relerr(r["trace"][0], z["losses"]) < 2 * TOL
    assert abs(r["scores"][0, 0] - float(z["score"])) < 2 * TOL * float(z["score"])


@pytest.mark.parametrize("tag", ["plain_", "conf_"])
def test_run_instance_fixture_with_shuffled_columns(engine, tag):
    """The reference's run_CGNN_tf / run_CGNN_confounders_tf on a DataFrame with shuffled columns, through the
    host layer's plan and the graph kernel."""
    import cgnn_b200.utils.Graph as my_graph
    from test_golden_reference import run_instance_case
    z = np.load(os.path.join(GOLD, "ref_shim_run_instance.npz"))
    plan, data, noise, theta0, h, tr, te = run_instance_case(z, tag, my_graph)
    r = engine.run_graph(data.astype(np.float32), [plan], n_runs=1, noise=noise[None].astype(np.float32),
                         init_params=theta0[None].astype(np.float32), want_trace=True, h_layer_dim=h,
                         train_epochs=tr, test_epochs=te, trace_every=1)
    assert relerr(r["trace"][0], z[tag + "losses"]) < 2 * TOL
    assert abs(r["scores"][0, 0] - float(z[tag + "score"])) < 2 * TOL * float(z[tag + "score"])


def test_graph_fourier_fixture(engine):
    """The graph kernel in random-feature mode against the reference's CGNN_tf(use_Fast_MMD=True): the training
    losses with the rp() matrices the reference drew (2e-4: the fp32 floor of this loss, tests/test_pairwise_gpu.py)."""
    import cgnn_b200.utils.Graph as my_graph
    from cgnn_b200.plan import build_plan
    z = np.load(os.path.join(GOLD, "ref_shim_graph_fourier.npz"))
    plan = build_plan(graph_from_fixture(z, my_graph))
    h, tr, k = int(z["h"]), int(z["train"]), int(z["k"])
    got = engine.graph_step_debug(z["data"], plan, z["noise"][:tr].astype(np.float32), z["theta0"].astype(np.float32),
                                  fourier_wz=z["wz"][:tr].astype(np.float32), h_layer_dim=h, use_Fast_MMD=True,
                                  nb_vectors_approx_MMD=k)
    assert relerr(got["loss"], z["losses"][:tr]) < 2e-4


def test_mmd_fixture_through_the_graph_kernel(engine):
    """Loss.MMD_loss_tf on raw matrices: a graph with no edges and zero input weights generates
    pred = b_out per column, so instead the fixture is checked by making `pred` the generator's
    output of a one-step instance: loss(step 0) is the MMD between data and the generated sample."""
    import cgnn_b200.utils.Graph as my_graph
    from cgnn_b200.plan import build_plan, n_params
    from oracle import cgnn as oc
    z = np.load(os.path.join(GOLD, "ref_shim_mmd.npz"))
    t = z["t1"].astype(np.float32)                      # 37 x 5
    g = my_graph.DirectedGraph()
    for k in range(4):
        g.add("V%d" % k, "V%d" % (k + 1), 1.0)
    plan = build_plan(g)
    rng = np.random.default_rng(0)
    theta = (0.3 * rng.normal(size=n_params(plan, 4))).astype(np.float32)
    noise = rng.normal(size=(1, 5, 37)).astype(np.float32)
    got = engine.graph_step_debug(t, plan, noise, theta, h_layer_dim=4, exp_policy=0)
    pred = got["pred"][0].astype(np.float64)
    # the reference's MMD_loss_tf semantics are pinned by test_golden_reference; here: kernel == that function
    from oracle import mmd
    want = mmd.mmd_loss_blocks(t.astype(np.float64), pred)
    assert abs(got["loss"][0] - want) < TOL * want
