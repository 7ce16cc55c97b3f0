"""Generates tests/golden/ref_shim_*.npz by executing the reference's own, unmodified model code
(/root/reference/Code/cgnn: utils/Loss.py, GNN.py, CGNN.py, CGNN_confounders.py) over
oracle.tf1_shim, in fp64 (TEST INFRASTRUCTURE ONLY; runs in the build container only).

What each fixture holds: the inputs (data, graph, initial parameters and every noise draw, in the
engine's layouts) and what the reference code computed from them (loss at every sess.run, the mean
returned by .evaluate(), MMD gradients).  tests/test_golden_reference.py replays them through the
oracle (CPU) and tests/test_golden_gpu.py through the CUDA kernels.

    python -m oracle.make_golden
"""
import contextlib
import io
import os
import sys

import numpy as np
import pandas as pd
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
OUT = os.path.join(ROOT, "tests", "golden")


def _quiet(fn, *a, **k):
    with contextlib.redirect_stdout(io.StringIO()):
        return fn(*a, **k)


def _split_draws(tf, n_init_tensors):
    init = [d[2] for d in tf.DRAWS if d[1] == "init"]
    runs = [d for d in tf.DRAWS if d[1] == "run"]
    assert len(init) == n_init_tensors
    return init, runs


def golden_mmd(tf, ref):
    from oracle import GAMMAS
    rng = np.random.default_rng(11)
    out = {}
    for k, (n, d) in enumerate(((50, 2), (37, 5), (64, 22))):
        t = rng.normal(size=(n, d))
        p = t[rng.permutation(n)] * 0.9 + 0.3 * rng.normal(size=(n, d))
        pv = tf.Variable(tf.constant(p))
        loss = ref["Loss"].MMD_loss_tf(tf.constant(t), pv)             # the reference's function, verbatim
        sess = tf.Session()
        sess.run(tf.global_variables_initializer())
        val = sess.run(loss)
        ctx_loss = loss.value(tf._Ctx(None, "run"))
        (g,) = torch.autograd.grad(ctx_loss, [pv.tensor])
        out.update({"t%d" % k: t, "p%d" % k: p, "loss%d" % k: val, "grad%d" % k: g.numpy()})
        tf.reset_default_graph()
    # the random-feature loss (Loss.py:35-56): record the matrix the reference's rp() drew
    for k, (n, d, nv) in enumerate(((40, 2, 6), (33, 5, 4))):
        t = rng.normal(size=(n, d))
        p = t[rng.permutation(n)] * 0.9 + 0.3 * rng.normal(size=(n, d))
        tf.reset(seed=40 + k)
        pv = tf.Variable(tf.constant(p))
        loss = ref["Loss"].Fourier_MMD_Loss_tf(tf.constant(t), pv, nv)
        sess = tf.Session()
        sess.run(tf.global_variables_initializer())
        val = sess.run(loss)
        draws = [d_[2] for d_ in tf.DRAWS if d_[1] == "run"]
        wz = np.concatenate([np.concatenate([2.0 * g * dr for g, dr in zip(GAMMAS, draws[:7])], 0), draws[7]], 1).T
        out.update({"ft%d" % k: t, "fp%d" % k: p, "fwz%d" % k: wz, "fk%d" % k: nv, "floss%d" % k: val})
        tf.reset_default_graph()
    np.savez_compressed(os.path.join(OUT, "ref_shim_mmd.npz"), **out)


def golden_pairwise(tf, ref):
    """GNN_tf.__init__/train/evaluate (GNN.py:33-126) and tf_run_instance (:135-166) on a bundled pair."""
    from cgnn_b200.utils import CCEPC_PairsFileReader
    df = CCEPC_PairsFileReader(os.path.join(ROOT, "examples", "Example_pairwise_pairs.csv"))
    out = {}
    for tag, (n, h, tr, te) in (("a", (96, 7, 14, 5)), ("b", (257, 30, 8, 3))):
        data = np.stack([df["A"][0][:n], df["B"][0][:n]], 1).astype("float32")
        tf.reset(seed=5 if tag == "a" else 6)
        model = ref["GNN"].GNN_tf(n, h_layer_dim=h, learning_rate=0.01, init_std=0.05, use_Fast_MMD=False)
        _quiet(model.train, data, verbose=False, train_epochs=tr)
        score = _quiet(model.evaluate, data, verbose=False, test_epochs=te)
        init, runs = _split_draws(tf, 4)
        W_in, b_in, W_out, b_out = init
        theta0 = np.concatenate([W_in[0], W_in[1], b_in, W_out[:, 0], b_out])
        noise = np.stack([d[2][:, 0] for d in runs])
        losses = np.array([r[0] for r in tf.RUN_LOG if len(r)], dtype=np.float64)
        assert noise.shape == (tr + te, n) and losses.shape == (tr + te,)
        out.update({tag + "_data": data, tag + "_h": h, tag + "_train": tr, tag + "_test": te, tag + "_theta0": theta0,
                    tag + "_noise": noise, tag + "_losses": losses, tag + "_score": score})
    # the same generator trained against the random-feature loss (use_Fast_MMD=True, GNN.py:69-70)
    from oracle import GAMMAS
    n, h, tr, te, nv = 72, 6, 9, 4, 5
    data = np.stack([df["A"][2][:n], df["B"][2][:n]], 1).astype("float32")
    tf.reset(seed=12)
    model = ref["GNN"].GNN_tf(n, h_layer_dim=h, learning_rate=0.01, init_std=0.05, use_Fast_MMD=True, nb_vectors_approx_MMD=nv)
    _quiet(model.train, data, verbose=False, train_epochs=tr)
    score = _quiet(model.evaluate, data, verbose=False, test_epochs=te)
    init, runs = _split_draws(tf, 4)
    W_in, b_in, W_out, b_out = init
    ids = sorted(set(d[0] for d in runs))                 # creation order: e, then rp(): 7 normal blocks, the phases
    per = {i: [d[2] for d in runs if d[0] == i] for i in ids}
    noise = np.stack([a[:, 0] for a in per[ids[0]]])
    wz = np.stack([np.concatenate([np.concatenate([2.0 * g * per[ids[1 + j]][s] for j, g in enumerate(GAMMAS)], 0),
                                   per[ids[8]][s]], 1).T for s in range(tr + te)])
    losses = np.array([r[0] for r in tf.RUN_LOG if len(r)], dtype=np.float64)
    out.update(dict(f_data=data, f_h=h, f_train=tr, f_test=te, f_k=nv, f_noise=noise, f_wz=wz, f_losses=losses, f_score=score,
                    f_theta0=np.concatenate([W_in[0], W_in[1], b_in, W_out[:, 0], b_out])))
    # both directions through tf_run_instance (column swap of GNN.py:162)
    n, h, tr, te = 64, 5, 6, 3
    data = np.stack([df["A"][1][:n], df["B"][1][:n]], 1).astype("float32")
    tf.reset(seed=9)
    ref["Settings"].SETTINGS.GPU = False
    xy, yx = _quiet(ref["GNN"].tf_run_instance, data, 0, 0, gpu=False, h_layer_dim=h, train_epochs=tr, test_epochs=te,
                    verbose=False)
    init = [d[2] for d in tf.DRAWS if d[1] == "init"]
    runs = [d[2][:, 0] for d in tf.DRAWS if d[1] == "run"]
    th = [np.concatenate([i[0][0], i[0][1], i[1], i[2][:, 0], i[3]]) for i in (init[:4], init[4:])]
    out.update(dict(c_data=data, c_h=h, c_train=tr, c_test=te, c_theta0=np.stack(th),
                    c_noise=np.stack(runs).reshape(2, tr + te, n), c_scores=np.array([xy, yx])))
    np.savez_compressed(os.path.join(OUT, "ref_shim_pairwise.npz"), **out)


def _small_graph(ref, conf):
    G = ref["Graph"]
    edges = [("V1", "V0"), ("V1", "V3"), ("V0", "V2"), ("V3", "V2"), ("V2", "V4"), ("V5", "V4")]
    skel = None
    if conf:
        skel = G.UndirectedGraph(pd.DataFrame(edges + [("V0", "V5")], columns=["Cause", "Effect"]))
    g = G.DirectedGraph(skeleton=skel) if conf else G.DirectedGraph()
    for k, (a, b) in enumerate(edges):
        g.add(a, b, 0.1 * (k + 1))
    return g


def golden_graph(tf, ref, conf):
    """CGNN_tf / CGNN_confounders_tf .__init__/train/evaluate (CGNN.py:35-153, CGNN_confounders.py:37-172)."""
    from oracle import cgnn as oc
    g = _small_graph(ref, conf)
    plan = oc.build_plan(g, confounders=conf)
    n, h, tr, te = 80, 6, 10, 4
    raw = pd.read_csv(os.path.join(ROOT, "examples", "Example_graph_numdata.csv"))[plan["list_nodes"]].values[:n]
    data = ((raw - raw.mean(0)) / raw.std(0)).astype("float32")
    tf.reset(seed=21 + conf)
    cls = ref["CGNN_confounders"].CGNN_confounders_tf if conf else ref["CGNN"].CGNN_tf
    model = cls(n, g, 0, 0, h_layer_dim=h, learning_rate=0.01, init_std=0.05, use_Fast_MMD=False)
    _quiet(model.train, data, verbose=False, train_epochs=tr)
    score = _quiet(model.evaluate, data, verbose=False, test_epochs=te)
    n_var = plan["n_var"]
    init, runs = _split_draws(tf, 4 * n_var)
    theta0 = np.concatenate([np.concatenate([t.reshape(-1) for t in init[4 * k:4 * k + 4]]) for k in range(n_var)])
    # noise draws of one sess.run in node-creation order: (confounders) one per skeleton edge, then the own
    # noise of every node in build order -> the engine's noise-column ids
    n_shared = plan["n_noise"] - n_var
    ids = sorted(set(d[0] for d in runs))
    assert len(ids) == plan["n_noise"]
    col_of = {}
    for pos, node_id in enumerate(ids):
        col_of[node_id] = (n_var + pos) if pos < n_shared else plan["nodes"][pos - n_shared]["col"]
    steps = tr + te
    noise = np.zeros((steps, plan["n_noise"], n))
    counter = {}
    for node_id, _, arr in runs:
        s = counter.get(node_id, 0)
        noise[s, col_of[node_id]] = arr[:, 0]
        counter[node_id] = s + 1
    losses = np.array([r[0] for r in tf.RUN_LOG if len(r)], dtype=np.float64)
    assert losses.shape == (steps,) and all(v == steps for v in counter.values())
    name = "confounders" if conf else "graph"
    np.savez_compressed(os.path.join(OUT, "ref_shim_%s.npz" % name), data=data, h=h, train=tr, test=te, theta0=theta0,
                        noise=noise, losses=losses, score=score,
                        edges=np.array(g.get_list_edges(order_by_weight=False, return_weights=False)),
                        weights=np.array([w for _, _, w in g.get_list_edges(order_by_weight=False)]),
                        skeleton=np.array(g.skeleton.get_list_edges_without_duplicate() if conf else []),
                        list_nodes=np.array(plan["list_nodes"]))


def golden_graph_fourier(tf, ref):
    """CGNN_tf with use_Fast_MMD=True (CGNN.py:92-93 -> Loss.py:35-56): the graph generator trained against the
    random-feature loss; the fixture keeps every matrix the reference's rp() drew."""
    from oracle import GAMMAS
    from oracle import cgnn as oc
    g = _small_graph(ref, False)
    plan = oc.build_plan(g)
    n, h, tr, te, nv = 64, 5, 7, 3, 4
    raw = pd.read_csv(os.path.join(ROOT, "examples", "Example_graph_numdata.csv"))[plan["list_nodes"]].values[:n]
    data = ((raw - raw.mean(0)) / raw.std(0)).astype("float32")
    tf.reset(seed=33)
    model = ref["CGNN"].CGNN_tf(n, g, 0, 0, h_layer_dim=h, learning_rate=0.01, init_std=0.05, use_Fast_MMD=True,
                                nb_vectors_approx_MMD=nv)
    _quiet(model.train, data, verbose=False, train_epochs=tr)
    score = _quiet(model.evaluate, data, verbose=False, test_epochs=te)
    n_var = plan["n_var"]
    init, runs = _split_draws(tf, 4 * n_var)
    theta0 = np.concatenate([np.concatenate([t.reshape(-1) for t in init[4 * k:4 * k + 4]]) for k in range(n_var)])
    ids = sorted(set(d[0] for d in runs))          # creation order: the nodes' noise (build order), then rp(): 7 normal blocks, phases
    assert len(ids) == n_var + 8
    per = {i: [d[2] for d in runs if d[0] == i] for i in ids}
    steps = tr + te
    noise = np.zeros((steps, plan["n_noise"], n))
    for pos in range(n_var):
        for s in range(steps):
            noise[s, plan["nodes"][pos]["col"]] = per[ids[pos]][s][:, 0]
    wz = np.stack([np.concatenate([np.concatenate([2.0 * gam * per[ids[n_var + j]][s] for j, gam in enumerate(GAMMAS)], 0),
                                   per[ids[n_var + 7]][s]], 1).T for s in range(steps)])
    assert wz.shape == (steps, n_var + 1, 7 * nv)
    losses = np.array([r[0] for r in tf.RUN_LOG if len(r)], dtype=np.float64)
    assert losses.shape == (steps,)
    np.savez_compressed(os.path.join(OUT, "ref_shim_graph_fourier.npz"), data=data, h=h, train=tr, test=te, k=nv, theta0=theta0,
                        noise=noise, wz=wz, losses=losses, score=score,
                        edges=np.array(g.get_list_edges(order_by_weight=False, return_weights=False)),
                        weights=np.array([w for _, _, w in g.get_list_edges(order_by_weight=False)]),
                        skeleton=np.array([]), list_nodes=np.array(plan["list_nodes"]))


def golden_run_instance(tf, ref):
    """run_CGNN_tf / run_CGNN_confounders_tf (CGNN.py:163-195, CGNN_confounders.py:182-216): the per-instance entry
    the search drivers call, fed with a DataFrame whose columns are NOT in the graph's node order - the reference
    re-orders them itself (`df_data[list_nodes]`).  Noise and initial parameters are stored per build position,
    not in any engine layout, so the replaying side has to do its own column mapping."""
    from oracle import cgnn as oc
    out = {}
    for conf in (False, True):
        g = _small_graph(ref, conf)
        plan = oc.build_plan(g, confounders=conf)
        n, h, tr, te = 48, 4, 6, 3
        raw = pd.read_csv(os.path.join(ROOT, "examples", "Example_graph_numdata.csv"))[plan["list_nodes"]].values[:n]
        std = ((raw - raw.mean(0)) / raw.std(0)).astype("float32")
        cols = list(plan["list_nodes"])
        shuffled = [cols[k] for k in np.random.default_rng(5 + conf).permutation(len(cols))]
        df = pd.DataFrame(std, columns=cols)[shuffled]
        tf.reset(seed=51 + conf)
        ref["Settings"].SETTINGS.GPU = True
        fn = ref["CGNN_confounders"].run_CGNN_confounders_tf if conf else ref["CGNN"].run_CGNN_tf
        score = _quiet(fn, df, g, 0, 0, gpu=True, nb_gpu=1, gpu_offset=0, h_layer_dim=h, train_epochs=tr, test_epochs=te,
                       verbose=False)
        n_var = plan["n_var"]
        init, runs = _split_draws(tf, 4 * n_var)
        theta0 = np.concatenate([np.concatenate([t.reshape(-1) for t in init[4 * k:4 * k + 4]]) for k in range(n_var)])
        ids = sorted(set(d[0] for d in runs))      # creation order: shared (one per skeleton edge), then own noise in build order
        per = {i: [d[2][:, 0] for d in runs if d[0] == i] for i in ids}
        n_shared = plan["n_noise"] - n_var
        steps = tr + te
        shared = np.stack([[per[ids[e]][s] for e in range(n_shared)] for s in range(steps)]) if n_shared else np.zeros((steps, 0, n))
        own = np.stack([[per[ids[n_shared + pos]][s] for pos in range(n_var)] for s in range(steps)])
        losses = np.array([r[0] for r in tf.RUN_LOG if len(r)], dtype=np.float64)
        assert losses.shape == (steps,)
        tag = "conf_" if conf else "plain_"
        out.update({tag + "data": df.values, tag + "columns": np.array(shuffled), tag + "h": h, tag + "train": tr, tag + "test": te,
                    tag + "theta0": theta0, tag + "own_noise": own, tag + "shared_noise": shared, tag + "losses": losses,
                    tag + "score": score,
                    tag + "edges": np.array(g.get_list_edges(order_by_weight=False, return_weights=False)),
                    tag + "weights": np.array([w for _, _, w in g.get_list_edges(order_by_weight=False)]),
                    tag + "skeleton": np.array(g.skeleton.get_list_edges_without_duplicate() if conf else [])})
    np.savez_compressed(os.path.join(OUT, "ref_shim_run_instance.npz"), **out)


def main():
    import refload
    from oracle import tf1_shim as tf
    assert refload.available(), "the reference tree is needed to generate the fixtures"
    tf.reset(0)
    ref = refload.load(tf.install())
    os.makedirs(OUT, exist_ok=True)
    golden_mmd(tf, ref)
    golden_pairwise(tf, ref)
    golden_graph(tf, ref, False)
    golden_graph(tf, ref, True)
    golden_graph_fourier(tf, ref)
    golden_run_instance(tf, ref)
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main()
