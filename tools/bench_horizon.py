"""Hill climbing on the bundled 22-variable example (reference settings) against the speculation horizon: candidates
scored per batch, number of batches, wasted candidates and wall time.  Usage: python tools/bench_horizon.py [h1 h2 ...]"""
import os, sys, time, copy
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pandas as pd
import cgnn
from cgnn_b200.CGNN import hill_climbing

EX = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "examples")
S = cgnn.SETTINGS
S.verbose = False
S.NB_RUNS, S.train_epochs, S.test_epochs, S.h_layer_dim = 32, 1000, 500, 20
data = pd.read_csv(os.path.join(EX, "Example_graph_numdata.csv"))
umg = cgnn.UndirectedGraph(pd.read_csv(os.path.join(EX, "Example_graph_skeleton.csv")))
dag0 = cgnn.GNN(backend="TensorFlow").orient_graph(data, umg)
ref = None
for h in [int(x) for x in sys.argv[1:]] or [1, 4, 9, 13, 18]:
    t0 = time.perf_counter()
    out = cgnn.CGNN(backend="TensorFlow").orient_directed_graph(data, copy.deepcopy(dag0), speculation_horizon=h)
    dt = time.perf_counter() - t0
    edges = sorted(out.get_list_edges(order_by_weight=False, return_weights=False))
    ref = ref or edges
    print("horizon %2d: %5.2f s, %2d batches %s = %d candidates, kernels %.2f s, same graph as first: %s"
          % (h, dt, len(hill_climbing.last_batches), hill_climbing.last_batches, sum(hill_climbing.last_batches),
             hill_climbing.last_kernel_seconds, edges == ref), flush=True)
