"""Tiny end-to-end pass over every kernel path, for compute-sanitizer (memcheck / racecheck): pairwise
(Gaussian and random-feature loss, ragged batch), graph (plain, hidden variables, cluster of 2), the
allocation-free entry point.  Usage: compute-sanitizer --tool memcheck python tools/sanitizer_target.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import torch
from util import make_pair
from test_graph_gpu import random_dag, make_data
from cgnn_b200 import Engine
from cgnn_b200.plan import build_plan
from cgnn_b200.noalloc import run_pairwise_noalloc

eng = Engine(0)
xy = np.zeros((2, 70, 2), np.float32)
xy[0] = make_pair(70, 1)
xy[1, :45] = make_pair(45, 2)
kw = dict(n_runs=2, h_layer_dim=7, train_epochs=3, test_epochs=2, seed=1)
r = eng.run_pairwise(xy, n_rows=np.array([70, 45], np.int32), exp_policy=3, want_trace=True, trace_every=1, **kw)
print("pairwise", r["scores"].ravel())
r = eng.run_pairwise(xy, n_rows=np.array([70, 45], np.int32), use_Fast_MMD=True, nb_vectors_approx_MMD=6, **kw)
print("fourier", r["scores"].ravel())
d = torch.from_numpy(xy[:1].copy()).cuda()
o = torch.zeros((1, 2, 2), device="cuda")
run_pairwise_noalloc(d, o, n_runs=2, h_layer_dim=7, train_epochs=2, test_epochs=1)
print("noalloc", o.cpu().numpy().ravel())
for conf in (False, True):
    g = random_dag(5, 2, 3, with_skeleton=conf)
    plan = build_plan(g, confounders=conf)
    data = make_data(70, 5, 1)
    for c in (1, 2):
        r = eng.run_graph(data, [plan, plan], n_runs=2, h_layer_dim=6, train_epochs=3, test_epochs=2, graph_cluster=c, seed=2)
        print("graph conf=%s cluster=%d" % (conf, c), r["scores"].ravel())
print("done")
