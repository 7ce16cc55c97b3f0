"""Parity of the CUDA graph path (CGNN and the hidden-variable variant) with the fp64 oracle,
through the C ABI.  Tolerance as in test_pairwise_gpu.py (north_star: 1e-5 relative)."""
import numpy as np
import pytest

from util import relerr

pytestmark = pytest.mark.gpu
TOL = 1e-5


def random_dag(n_var, max_par, seed, with_skeleton=False):
    """A DirectedGraph over V0..V{n-1} (random order of edge insertion) and matching data columns."""
    from cgnn_b200.utils.Graph import DirectedGraph, UndirectedGraph
    import pandas as pd
    rng = np.random.default_rng(seed)
    perm = rng.permutation(n_var)
    edges = []
    for k in range(1, n_var):
        for p in rng.choice(k, size=min(k, int(rng.integers(1, max_par + 1))), replace=False):
            edges.append(("V%d" % perm[p], "V%d" % perm[k]))
    order = rng.permutation(len(edges))
    skel = None
    if with_skeleton:
        extra = [("V%d" % perm[0], "V%d" % perm[n_var - 1])] if n_var > 3 else []
        skel = UndirectedGraph(pd.DataFrame([edges[o] for o in order] + extra, columns=["Cause", "Effect"]))
    g = DirectedGraph(skeleton=skel) if with_skeleton else DirectedGraph()
    for o in order:
        g.add(edges[o][0], edges[o][1], float(rng.uniform(0.1, 1)))
    return g


def make_data(n, n_var, seed):
    rng = np.random.default_rng(seed)
    z = rng.normal(size=(n, n_var))
    mix = rng.normal(size=(n_var, n_var)) * 0.5 + np.eye(n_var)
    d = np.tanh(z @ mix) + 0.1 * z
    if n == 1:
        return d.astype(np.float32)
    return ((d - d.mean(0)) / d.std(0)).astype(np.float32)


def _oracle_steps(data, plan, noise, theta0, h):
    from oracle import cgnn as oc
    from oracle.adam import TFAdam
    theta = theta0.astype(np.float64)
    opt = TFAdam(theta.size, 0.01)
    out = dict(loss=[], pred=[], dldp=[], grad=[], params=[])
    for it in range(noise.shape[0]):
        r = oc.step_debug(theta, data.astype(np.float64), noise[it].astype(np.float64), plan, h)
        theta = opt.step(theta, r["grad"])
        out["loss"].append(r["loss"]); out["pred"].append(r["pred"]); out["dldp"].append(r["dloss_dpred"])
        out["grad"].append(r["grad"]); out["params"].append(theta.copy())
    return {k: np.array(v) for k, v in out.items()}


@pytest.mark.parametrize("pol", [0, 1, 3])          # 1 is what the graph path runs by default (CGNN_EXP_POLICY_GRAPH)
@pytest.mark.parametrize("n,n_var,h,conf", [(100, 5, 8, False), (257, 11, 20, False), (64, 22, 20, False),
                                            (33, 3, 32, False), (96, 8, 12, True), (130, 19, 20, True), (1, 2, 4, False),
                                            (70, 21, 20, False), (45, 23, 24, False), (40, 17, 9, True)])
def test_graph_step_debug_matches_oracle(engine, n, n_var, h, conf, pol):
    from cgnn_b200.plan import build_plan, n_params
    from oracle import cgnn as oc
    g = random_dag(n_var, 3, 7 * n_var + n, with_skeleton=conf)
    plan = build_plan(g, confounders=conf)
    oplan = oc.build_plan(g, confounders=conf)
    assert [(a["col"], a["parents"], a["noise"]) for a in plan["nodes"]] == \
           [(b["col"], b["parents"], b["noise"]) for b in oplan["nodes"]]
    data = make_data(n, n_var, n)
    rng = np.random.default_rng(n + n_var)
    P = n_params(plan, h)
    assert P == oc.n_params(oplan, h)
    theta0 = (0.05 * rng.normal(size=P)).astype(np.float32)
    steps = 3
    noise = rng.normal(size=(steps, plan["n_noise"], n)).astype(np.float32)
    got = engine.graph_step_debug(data, plan, noise, theta0, h_layer_dim=h, exp_policy=pol)
    ref = _oracle_steps(data, oplan, noise, theta0, h)
    for s in range(steps):
        assert abs(got["loss"][s] - ref["loss"][s]) <= TOL * abs(ref["loss"][s]), (s, got["loss"][s], ref["loss"][s])
        assert relerr(got["pred"][s], ref["pred"][s]) < TOL
        assert relerr(got["dldp"][s], ref["dldp"][s]) < TOL
        assert relerr(got["grad"][s], ref["grad"][s]) < TOL
        # Adam divides by sqrt(v)+1e-8: for parameters whose gradient is itself ~1e-8 a 1e-10 gradient
        # difference moves the update by ~lr*1e-2, so post-update parameters get a looser bound
        assert relerr(got["params"][s], ref["params"][s]) < 1e-4


@pytest.mark.parametrize("pol", [-1, 3])            # -1: the per-path default
@pytest.mark.parametrize("n,n_var,h", [(500, 22, 20), (1500, 20, 20)])
def test_graph_full_size_step(engine, n, n_var, h, pol):
    """The shapes of the bundled graph example (500 rows, 22 variables, h = 20: both matrices in shared memory) and
    of BASELINE configs[4] (1500 rows, 20 variables: the observed matrix streamed through the per-warp ring)."""
    from cgnn_b200.plan import build_plan, n_params
    from oracle import cgnn as oc
    g = random_dag(n_var, 2, 5)
    plan = build_plan(g)
    data = make_data(n, n_var, 3)
    rng = np.random.default_rng(1)
    theta0 = (0.05 * rng.normal(size=n_params(plan, h))).astype(np.float32)
    noise = rng.normal(size=(2, plan["n_noise"], n)).astype(np.float32)
    got = engine.graph_step_debug(data, plan, noise, theta0, h_layer_dim=h, exp_policy=pol)
    ref = _oracle_steps(data, oc.build_plan(g), noise, theta0, h)
    for s in range(2):
        assert abs(got["loss"][s] - ref["loss"][s]) <= TOL * abs(ref["loss"][s])
        assert relerr(got["dldp"][s], ref["dldp"][s]) < TOL
        assert relerr(got["grad"][s], ref["grad"][s]) < TOL


@pytest.mark.parametrize("pol", [0, -1])            # -1: the policy production runs
def test_run_graph_external_noise_matches_oracle(engine, pol):
    from cgnn_b200.plan import build_plan, n_params
    from oracle import cgnn as oc
    n, n_var, h, tr, te, runs = 90, 6, 10, 20, 5, 2
    graphs = [random_dag(n_var, 2, 40 + k) for k in range(3)]
    plans = [build_plan(g, columns=["V%d" % c for c in range(n_var)]) for g in graphs]
    data = make_data(n, n_var, 9)
    rng = np.random.default_rng(4)
    ld = max(n_params(p, h) for p in plans)
    n_inst = len(plans) * runs
    init = (0.05 * rng.normal(size=(n_inst, ld))).astype(np.float32)
    noise = rng.normal(size=(n_inst, tr + te, n_var, n)).astype(np.float32)
    r = engine.run_graph(data, plans, n_runs=runs, noise=noise, init_params=init, want_trace=True, want_params=True,
                         h_layer_dim=h, train_epochs=tr, test_epochs=te, trace_every=1, exp_policy=pol)
    for gi, plan in enumerate(plans):
        P = n_params(plan, h)
        for run in range(runs):
            k = gi * runs + run
            theta, losses, _ = oc.train(init[k, :P], data.astype(np.float64), noise[k, :tr].astype(np.float64), plan, h)
            score, ev = oc.evaluate(theta, data.astype(np.float64), noise[k, tr:].astype(np.float64), plan, h)
            assert relerr(r["trace"][k, :tr], losses) < 2e-5
            assert relerr(r["trace"][k, tr:], ev) < 2e-5
            assert relerr(r["params"][k, :P], theta) < 1e-4
            assert abs(r["scores"][gi, run] - score) < 2e-5 * abs(score)


def test_run_graph_determinism_subsets_and_limits(engine):
    from cgnn_b200 import EngineError
    from cgnn_b200.plan import build_plan
    n_var = 7
    g = random_dag(n_var, 2, 3)
    cols = ["V%d" % c for c in range(n_var)]
    plans = [build_plan(g, columns=cols)] * 2
    data = make_data(300, n_var, 2)
    kw = dict(n_runs=3, h_layer_dim=10, train_epochs=15, test_epochs=4, seed=77)
    a = engine.run_graph(data, plans, **kw)["scores"].copy()
    b = engine.run_graph(data, plans, **kw)["scores"].copy()
    assert np.array_equal(a, b) and np.all(np.isfinite(a)) and np.all(a > 0)
    assert not np.array_equal(a[0], a[1])            # same graph, different Philox streams
    # the second graph alone with the matching uid_base reproduces its rows
    c = engine.run_graph(data, plans[1:], uid_base=3, **kw)["scores"]
    assert np.array_equal(a[1:], c)
    # row subsets (CGNN.py:183-185): identity subset == no subset
    ident = np.tile(np.arange(300, dtype=np.int32), (6, 1))
    d = engine.run_graph(data, plans, row_subsets=ident, **kw)["scores"]
    assert np.array_equal(a, d)
    sub = np.stack([np.random.default_rng(k).permutation(300)[:200] for k in range(6)]).astype(np.int32)
    e = engine.run_graph(data, plans, row_subsets=sub, **kw)["scores"]
    one = engine.run_graph(np.ascontiguousarray(data[sub[4]]), plans[1:], n_runs=1, uid_base=4,
                           **{k: v for k, v in kw.items() if k != "n_runs"})["scores"]
    assert np.array_equal(e[1, 1], one[0, 0])
    with pytest.raises(EngineError):
        engine.run_graph(data, plans, n_runs=1, h_layer_dim=33)
    with pytest.raises(EngineError):
        engine.run_graph(np.zeros((1600, n_var), np.float32), plans, n_runs=1)


def test_generate_matches_oracle_and_fits(engine):
    """CGNN_tf.generate (CGNN.py:155-160): forward sample of a fitted model."""
    import pandas as pd
    import cgnn_b200 as cgnn
    from cgnn_b200 import runtime
    from cgnn_b200.plan import build_plan, n_params
    from oracle import cgnn as oc
    n_var, h, n = 5, 8, 150
    g = random_dag(n_var, 2, 12)
    plan = build_plan(g)
    rng = np.random.default_rng(0)
    theta = (0.3 * rng.normal(size=n_params(plan, h))).astype(np.float32)
    noise = rng.normal(size=(plan["n_noise"], n)).astype(np.float32)
    got = engine.graph_generate(plan, theta, noise, h_layer_dim=h)
    want = oc.generate(theta.astype(np.float64), noise.astype(np.float64), oc.build_plan(g), h)
    assert relerr(got, want) < TOL
    # public API: fit then sample; a trained model's sample must be much closer (in MMD) to the data than an
    # untrained model's
    data = pd.DataFrame(make_data(300, n_var, 5), columns=plan["list_nodes"])
    try:
        sample = cgnn.CGNN().generate(data, g, h_layer_dim=10, train_epochs=300, test_epochs=1)
        raw = cgnn.CGNN().generate(data, g, h_layer_dim=10, train_epochs=0, test_epochs=1)
    finally:
        runtime.close_engines()
    assert list(sample.columns) == [c for c in data.columns] and sample.shape == data.shape
    from oracle import mmd
    d = data.values.astype(np.float64)
    d = (d - d.mean(0)) / d.std(0)
    assert mmd.mmd_loss_blocks(d, sample.values) < 0.3 * mmd.mmd_loss_blocks(d, raw.values)


def test_cluster_sizes_agree(engine):
    """One instance on a cluster of 2-4 CTAs (distributed shared memory gradient sum) gives the scores of the
    single-CTA run: the gradient accumulators are fixed point, so the parameter trajectories are identical."""
    from cgnn_b200.plan import build_plan
    n_var = 9
    g = random_dag(n_var, 2, 8)
    plan = build_plan(g)
    data = make_data(700, n_var, 4)          # 22 row blocks: uneven split over 3 and 4 CTAs
    kw = dict(n_runs=3, h_layer_dim=14, train_epochs=12, test_epochs=4, seed=9, want_params=True, want_trace=True, trace_every=3)
    base = engine.run_graph(data, [plan, plan], graph_cluster=1, **kw)
    assert engine.last_graph_cluster == 1
    for c in (2, 3, 4, 6, 8):
        r = engine.run_graph(data, [plan, plan], graph_cluster=c, **kw)
        assert engine.last_graph_cluster == c
        assert np.array_equal(r["params"], base["params"]), "cluster size %d changed the trained parameters" % c
        np.testing.assert_allclose(r["scores"], base["scores"], rtol=1e-6)
        np.testing.assert_allclose(r["trace"], base["trace"], rtol=1e-6, equal_nan=True)
    # 20 variables x 1500 rows: the observed matrix no longer fits shared memory and is streamed through the
    # per-warp cp.async ring - the same property must hold on that code path
    g20 = random_dag(20, 3, 5)
    plan20 = build_plan(g20)
    data20 = make_data(1500, 20, 6)
    kw20 = dict(n_runs=2, h_layer_dim=20, train_epochs=4, test_epochs=2, seed=3, want_params=True)
    base20 = engine.run_graph(data20, [plan20], graph_cluster=1, **kw20)
    for c in (2, 3):
        r = engine.run_graph(data20, [plan20], graph_cluster=c, **kw20)
        assert engine.last_graph_cluster == c
        assert np.array_equal(r["params"], base20["params"]), "cluster size %d changed the trained parameters (ring path)" % c
        np.testing.assert_allclose(r["scores"], base20["scores"], rtol=1e-6)
    # automatic choice: few long instances -> a cluster; many instances -> one CTA each
    engine.run_graph(make_data(1500, n_var, 1), [plan], n_runs=4, h_layer_dim=10, train_epochs=2, test_epochs=1)
    assert engine.last_graph_cluster >= 2
    engine.run_graph(make_data(512, n_var, 2), [plan] * 10, n_runs=32, h_layer_dim=10, train_epochs=1, test_epochs=1)
    assert engine.last_graph_cluster == 1


# ---- random-feature ("Fourier") MMD on the graph path (CGNN.py:92-93, Loss.py:35-56) ----------------------
FTOL = 2e-4        # fp32 floor of this loss, see tests/test_pairwise_gpu.py


@pytest.mark.parametrize("n,n_var,h,k,conf", [(100, 5, 8, 6, False), (500, 22, 20, 100, False), (130, 19, 20, 10, True), (33, 3, 32, 3, False)])
def test_graph_fourier_step_matches_oracle(engine, n, n_var, h, k, conf):
    from cgnn_b200.plan import build_plan, n_params
    from oracle import cgnn as oc
    from oracle.mmd import rp_literal
    g = random_dag(n_var, 3, 3 * n_var + n, with_skeleton=conf)
    plan = build_plan(g, confounders=conf)
    oplan = oc.build_plan(g, confounders=conf)
    data = make_data(n, n_var, n)
    rng = np.random.RandomState(n + n_var)
    theta0 = (0.05 * rng.standard_normal(n_params(plan, h))).astype(np.float32)
    steps = 2
    noise = rng.standard_normal((steps, plan["n_noise"], n)).astype(np.float32)
    wz = np.stack([rp_literal(k, n_var, rng) for _ in range(steps)]).astype(np.float32)
    got = engine.graph_step_debug(data, plan, noise, theta0, fourier_wz=wz, h_layer_dim=h, use_Fast_MMD=True, nb_vectors_approx_MMD=k)
    for s in range(steps):
        th = theta0 if s == 0 else got["params"][s - 1]
        r = oc.step_debug(th.astype(np.float64), data.astype(np.float64), noise[s].astype(np.float64), oplan, h,
                          wz=wz[s].astype(np.float64), k=k)
        assert abs(got["loss"][s] - r["loss"]) <= FTOL * abs(r["loss"]), (s, got["loss"][s], r["loss"])
        assert relerr(got["dldp"][s], r["dloss_dpred"]) < FTOL
        assert relerr(got["grad"][s], r["grad"]) < FTOL
        # ... and measured against the reference's own arithmetic: its literal graph in fp32 (CGNN.py:92-93) vs fp64
        import torch
        r32 = oc.step_debug(th.astype(np.float32), data.astype(np.float32), noise[s].astype(np.float32), oplan, h,
                            dtype=torch.float32, wz=wz[s].astype(np.float32), k=k)
        for mine, a32, a64 in ((got["dldp"][s], r32["dloss_dpred"], r["dloss_dpred"]), (got["grad"][s], r32["grad"], r["grad"])):
            assert relerr(mine, a64) <= max(2.0 * relerr(a32, a64), 2e-5)


def test_graph_fourier_runs_on_clusters(engine):
    """N = 1500 needs three CTAs per instance in this mode; cluster sizes must agree bit for bit."""
    from cgnn_b200.plan import build_plan
    n_var = 6
    plan = build_plan(random_dag(n_var, 2, 2))
    kw = dict(n_runs=2, h_layer_dim=10, train_epochs=6, test_epochs=3, seed=4, use_Fast_MMD=True, nb_vectors_approx_MMD=20,
              want_params=True)
    small = make_data(480, n_var, 3)
    a = engine.run_graph(small, [plan], graph_cluster=1, **kw)
    b = engine.run_graph(small, [plan], graph_cluster=3, **kw)
    assert np.array_equal(a["params"], b["params"]) and np.allclose(a["scores"], b["scores"], rtol=1e-6)
    big = engine.run_graph(make_data(1500, n_var, 5), [plan], **kw)
    assert engine.last_graph_cluster >= 3 and np.all(np.isfinite(big["scores"])) and np.all(big["scores"] > 0)


def test_graph_size_limits(engine):
    """Largest shapes: 1536 rows x 20 variables fits; combinations that exceed one SM's shared memory are
    refused with a message that says so (no silent truncation, no crash)."""
    from cgnn_b200 import EngineError
    from cgnn_b200.plan import build_plan
    plan20 = build_plan(random_dag(20, 2, 6))
    r = engine.run_graph(make_data(1536, 20, 1), [plan20], n_runs=2, h_layer_dim=20, train_epochs=2, test_epochs=1)
    assert np.all(np.isfinite(r["scores"]))
    plan32 = build_plan(random_dag(32, 3, 7))
    with pytest.raises(EngineError, match="shared memory"):
        engine.run_graph(make_data(1536, 32, 1), [plan32], n_runs=1, h_layer_dim=32, train_epochs=1, test_epochs=1)
    r = engine.run_graph(make_data(512, 32, 2), [plan32], n_runs=1, h_layer_dim=20, train_epochs=2, test_epochs=1)
    assert np.all(np.isfinite(r["scores"]))
    with pytest.raises(EngineError):
        engine.run_graph(make_data(64, 33, 2), [build_plan(random_dag(33, 2, 8))], n_runs=1)


@pytest.mark.parametrize("pol", [0, 1, 3])
def test_graph_trained_state_parity(engine, pol):
    """After 150 Adam steps the loss has dropped and dLoss/dPred is a small difference of large PT / PP sums:
    the case that stresses the fp32 accumulation scheme (re-anchored on the kernel's own parameters)."""
    from cgnn_b200.plan import build_plan, n_params
    from oracle import cgnn as oc
    n, n_var, h, steps = 256, 6, 12, 150
    g = random_dag(n_var, 2, 21)
    plan = build_plan(g)
    oplan = oc.build_plan(g)
    rng = np.random.default_rng(6)
    # data that the graph can actually fit: sampled from a random generator of the same structure
    th_gen = (0.8 * rng.normal(size=n_params(plan, h))).astype(np.float64)
    data = oc.generate(th_gen, rng.normal(size=(plan["n_noise"], n)), oplan, h)
    data = ((data - data.mean(0)) / data.std(0)).astype(np.float32)
    theta0 = (0.05 * rng.normal(size=n_params(plan, h))).astype(np.float32)
    noise = rng.normal(size=(steps, plan["n_noise"], n)).astype(np.float32)
    got = engine.graph_step_debug(data, plan, noise, theta0, h_layer_dim=h, exp_policy=pol)
    assert got["loss"][-1] < 0.5 * got["loss"][0]
    th = got["params"][steps - 2].astype(np.float64)
    r = oc.step_debug(th, data.astype(np.float64), noise[steps - 1].astype(np.float64), oplan, h)
    assert abs(got["loss"][-1] - r["loss"]) <= TOL * abs(r["loss"]), (got["loss"][-1], r["loss"])
    assert relerr(got["dldp"][-1], r["dloss_dpred"]) < TOL
    assert relerr(got["grad"][-1], r["grad"]) < TOL


@pytest.mark.parametrize("n,n_var,steps", [(500, 22, 160), (1500, 20, 120)])
def test_graph_trained_state_at_the_benchmark_shapes(engine, n, n_var, steps):
    """The same at the shapes BASELINE.json quotes (bundled example: 500 rows x 22 variables, both matrices in
    shared memory; configs[4]: 1500 rows x 20 variables, observed matrix streamed), default exp policy: after
    the loss has dropped, ONE fp64 oracle step re-anchored on the kernel's own parameters - loss, dLoss/dPred and
    the parameter gradient within 1e-5 (Loss.py:12-32 after CGNN.py:116-121 has run `steps` times)."""
    from cgnn_b200.plan import build_plan, n_params
    from oracle import cgnn as oc
    h = 20
    g = random_dag(n_var, 2, 31)
    plan = build_plan(g)
    oplan = oc.build_plan(g)
    rng = np.random.default_rng(16)
    th_gen = (0.8 * rng.normal(size=n_params(plan, h))).astype(np.float64)
    data = oc.generate(th_gen, rng.normal(size=(plan["n_noise"], n)), oplan, h)
    data = ((data - data.mean(0)) / data.std(0)).astype(np.float32)
    theta0 = (0.05 * rng.normal(size=n_params(plan, h))).astype(np.float32)
    noise = rng.normal(size=(steps, plan["n_noise"], n)).astype(np.float32)
    got = engine.graph_step_debug(data, plan, noise, theta0, h_layer_dim=h)
    assert got["loss"][-1] < 0.5 * got["loss"][0]
    th = got["params"][steps - 2].astype(np.float64)
    r = oc.step_debug(th, data.astype(np.float64), noise[steps - 1].astype(np.float64), oplan, h)
    assert abs(got["loss"][-1] - r["loss"]) <= TOL * abs(r["loss"]), (got["loss"][-1], r["loss"])
    assert relerr(got["dldp"][-1], r["dloss_dpred"]) < TOL
    assert relerr(got["grad"][-1], r["grad"]) < TOL


def _wide_confounder_graph(n_var=20, n_skel=30, seed=3):
    """An oriented 20-variable graph over a 30-edge skeleton (max_nb_points-sized problems of the hidden-variable model)."""
    import pandas as pd
    from cgnn_b200.utils.Graph import DirectedGraph, UndirectedGraph
    rng = np.random.default_rng(seed)
    pairs = set()
    for k in range(1, n_var):                                   # a spanning tree first: every variable has a neighbour
        pairs.add((int(rng.integers(0, k)), k))
    while len(pairs) < n_skel:
        a, b = sorted(rng.choice(n_var, 2, replace=False).tolist())
        pairs.add((a, b))
    pairs = sorted(pairs)
    skel = UndirectedGraph(pd.DataFrame([("V%d" % a, "V%d" % b) for a, b in pairs], columns=["Cause", "Effect"]))
    g = DirectedGraph(skeleton=skel)
    for a, b in pairs:
        g.add("V%d" % a, "V%d" % b, float(rng.uniform(0.1, 1)))   # low index -> high index: acyclic
    return g


def test_wide_hidden_variable_model_at_1500_rows(engine):
    """20 variables, a 30-edge skeleton (50 noise columns), SETTINGS.max_nb_points = 1500 rows: buffering every noise column
    per row block would need more shared memory than an SM has, so the engine re-derives the noise at every use
    (CGNN_confounders.py:66-96 at the reference's default size).  One train step against the fp64 oracle, then a short
    scored batch."""
    from cgnn_b200.plan import build_plan, n_params
    from oracle import cgnn as oc
    n, h = 1500, 20
    g = _wide_confounder_graph()
    plan = build_plan(g, confounders=True)
    oplan = oc.build_plan(g, confounders=True)
    assert plan["n_noise"] == 50
    data = make_data(n, 20, 8)
    rng = np.random.default_rng(2)
    theta0 = (0.05 * rng.normal(size=n_params(plan, h))).astype(np.float32)
    noise = rng.normal(size=(1, plan["n_noise"], n)).astype(np.float32)
    got = engine.graph_step_debug(data, plan, noise, theta0, h_layer_dim=h)
    assert engine.last_graph_mode & 4, "expected the noise-regeneration plan at this size"
    ref = _oracle_steps(data, oplan, noise, theta0, h)
    assert abs(got["loss"][0] - ref["loss"][0]) <= TOL * abs(ref["loss"][0])
    assert relerr(got["dldp"][0], ref["dldp"][0]) < TOL
    assert relerr(got["grad"][0], ref["grad"][0]) < TOL
    r = engine.run_graph(data, [plan], n_runs=3, h_layer_dim=h, train_epochs=6, test_epochs=3, seed=5)
    assert engine.last_graph_mode & 4
    assert np.all(np.isfinite(r["scores"])) and np.all(r["scores"] > 0) and len(set(r["scores"].ravel().tolist())) == 3


@pytest.mark.parametrize("conf", [False, True])
def test_noise_regeneration_is_bit_identical(engine, conf):
    """The same instances with the noise buffered per row block and re-derived at every use (in-kernel Philox streams):
    identical trained parameters and scores."""
    from cgnn_b200.plan import build_plan
    g = random_dag(9, 2, 14, with_skeleton=conf)
    plan = build_plan(g, confounders=conf)
    data = make_data(300, 9, 4)
    kw = dict(n_runs=3, h_layer_dim=12, train_epochs=10, test_epochs=4, seed=21, want_params=True)
    a = engine.run_graph(data, [plan], **kw)
    assert not engine.last_graph_mode & 4
    b = engine.run_graph(data, [plan], graph_noise_regen=1, **kw)
    assert engine.last_graph_mode & 4
    assert np.array_equal(a["params"], b["params"]) and np.array_equal(a["scores"], b["scores"])
