"""Throughput of the random-feature (use_Fast_MMD) mode next to the Gaussian-kernel MMD, same shapes."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, pandas as pd
from cgnn_b200 import Engine, synth, DirectedGraph
from cgnn_b200.plan import build_plan
from cgnn_b200.utils.scaling import scale
eng = Engine(0)
xy, _ = synth.make_pairs(0, 37, 1500)
for fast in (False, True):
    for rep in range(2):
        r = eng.run_pairwise(xy, n_runs=4, h_layer_dim=30, train_epochs=100, test_epochs=50, seed=1, use_Fast_MMD=fast, nb_vectors_approx_MMD=100)
    print("pairwise N=1500 fast=%s: 296 instances x 150 steps: %.1f ms -> %.0f train-steps/s" % (fast, eng.last_kernel_ms, 296 * 100 / eng.last_kernel_ms * 1e3))
EX = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "examples")
data = pd.read_csv(os.path.join(EX, "Example_graph_numdata.csv"))
dag = DirectedGraph(pd.read_csv(os.path.join(EX, "Example_graph_target.csv")))
plan = build_plan(dag, columns=list(data.columns))
X = scale(data.values).astype(np.float32)
for fast in (False, True):
    for rep in range(2):
        r = eng.run_graph(X, [plan] * 4, n_runs=37, h_layer_dim=20, train_epochs=40, test_epochs=20, seed=1, use_Fast_MMD=fast, nb_vectors_approx_MMD=100)
    print("graph N=500 d=22 fast=%s: 148 instances x 60 steps: %.1f ms (cluster %d)" % (fast, eng.last_kernel_ms, eng.last_graph_cluster))
