"""The oracle against itself: the block / closed-form restatement used by the CUDA kernels equals
autograd of the literal Loss.py graph; TF-Adam; generator plans."""
import numpy as np
import pytest
import torch


def test_block_form_and_analytic_gradient_equal_literal_graph():
    from oracle import mmd
    rng = np.random.default_rng(0)
    for n, d in ((60, 2), (45, 7), (33, 22)):
        t = rng.normal(size=(n, d))
        p = t + 0.3 * rng.normal(size=(n, d))
        loss, grad = mmd.mmd_loss_and_grad_autograd(t, p)
        assert abs(mmd.mmd_loss_blocks(t, p) - loss) < 1e-10 * abs(loss)
        assert np.max(np.abs(mmd.mmd_grad_analytic(t, p) - grad)) < 1e-10 * np.max(np.abs(grad))
    # identical samples -> exactly zero MMD in exact arithmetic
    assert abs(mmd.mmd_loss_blocks(t, t)) < 1e-12


def test_literal_graph_fp32_noise_floor():
    """The reference's own arithmetic (fp32 expansion form) differs from fp64 by ~1e-7..1e-6: the 1e-5
    parity bound leaves an order of magnitude."""
    from oracle import mmd
    rng = np.random.default_rng(1)
    t = rng.normal(size=(300, 2)); p = t + 0.05 * rng.normal(size=(300, 2))
    l64 = float(mmd.mmd_loss_literal(torch.tensor(t), torch.tensor(p)))
    l32 = float(mmd.mmd_loss_literal(torch.tensor(t, dtype=torch.float32), torch.tensor(p, dtype=torch.float32)))
    assert abs(l32 - l64) < 5e-4 * abs(l64)


def test_tf_adam_first_steps():
    from oracle.adam import TFAdam
    opt = TFAdam(3, lr=0.01)
    th = np.array([1.0, -2.0, 0.5])
    g = np.array([0.3, -0.1, 0.0])
    th1 = opt.step(th, g)
    # first step: m = 0.1 g, v = 0.001 g^2, lr_t = lr*sqrt(1-b2)/(1-b1) -> update = lr * g/(|g| + eps*sqrt(1-b2)...)
    lr_t = 0.01 * np.sqrt(1 - 0.999) / (1 - 0.9)
    want = th - lr_t * (0.1 * g) / (np.sqrt(0.001 * g * g) + 1e-8)
    np.testing.assert_allclose(th1, want, rtol=1e-14)
    assert th1[2] == 0.5


def test_pairwise_training_reduces_loss():
    from oracle import gnn
    rng = np.random.default_rng(2)
    x = rng.uniform(-1, 1, 120); y = np.tanh(2 * x) + 0.2 * rng.normal(size=120)
    d = np.stack([(x - x.mean()) / x.std(), (y - y.mean()) / y.std()], 1)
    th0 = 0.05 * rng.normal(size=gnn.n_params(8))
    _, losses, _ = gnn.train(th0, d, rng.normal(size=(60, 120)), 8)
    assert losses[-1] < 0.3 * losses[0]


def test_graph_plan_and_forward_shapes():
    import sys, os
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from test_graph_gpu import random_dag, make_data
    from oracle import cgnn as oc
    from cgnn_b200.plan import build_plan, n_params
    for conf in (False, True):
        g = random_dag(9, 3, 4, with_skeleton=conf)
        op = oc.build_plan(g, confounders=conf)
        mp = build_plan(g, confounders=conf)
        assert [(a["col"], a["parents"], a["noise"]) for a in op["nodes"]] == [(b["col"], b["parents"], b["noise"]) for b in mp["nodes"]]
        P = oc.n_params(op, 6)
        assert P == n_params(mp, 6)
        rng = np.random.default_rng(0)
        r = oc.step_debug(0.05 * rng.normal(size=P), make_data(40, 9, 1), rng.normal(size=(op["n_noise"], 40)), op, 6)
        assert r["pred"].shape == (40, 9) and r["grad"].shape == (P,) and np.isfinite(r["loss"])


def test_far_tile_threshold_bound():
    """The pairwise kernel leaves the gamma = 50 kernel out for tile pairs whose cause ranges are more than
    CGNN_FAR_DP (primed units, D' = 0.25 log2(e) D) apart.  The dropped term against the smallest retained one,
    and against the whole weight sum of Loss.py:10's bandwidth list, at the threshold (it only shrinks beyond)."""
    import math
    import re
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    txt = open(os.path.join(root, "cgnn_b200", "csrc", "common.cuh")).read()
    far_dp = float(re.search(r"#define\s+CGNN_FAR_DP\s+([0-9.]+)f", txt).group(1))
    d = far_dp / (0.25 * math.log2(math.e))
    gam = [0.005, 0.05, 0.25, 0.5, 1.0, 5.0, 50.0]
    dropped_w, dropped_k = 50.0 * math.exp(-50.0 * d), math.exp(-50.0 * d)
    kept_w = sum(g * math.exp(-g * d) for g in gam[:-1])
    kept_k = sum(math.exp(-g * d) for g in gam[:-1])
    assert dropped_w / (0.005 * math.exp(-0.005 * d)) < 1e-8          # against the widest kernel alone
    assert dropped_w / kept_w < 1e-10 and dropped_k / kept_k < 1e-12   # against what is actually summed
    # second level ("very far"): the gamma = 5 kernel goes as well
    far2_dp = float(re.search(r"#define\s+CGNN_FAR2_DP\s+([0-9.]+)f", txt).group(1))
    d2 = far2_dp / (0.25 * math.log2(math.e))
    assert (5.0 * math.exp(-5.0 * d2) + 50.0 * math.exp(-50.0 * d2)) / (0.005 * math.exp(-0.005 * d2)) < 1e-8
    assert (math.exp(-5.0 * d2) + math.exp(-50.0 * d2)) / sum(math.exp(-g * d2) for g in gam[:-2]) < 1e-10
