"""Throughput of the graph instance kernel on the two graph configurations of BASELINE.json:
  C2: bundled 22-variable example (N=500, h=20, nb_runs=32): the start DAG + every acyclic single-edge
      reversal = one converged hill-climbing sweep (1+27 graphs x 32 runs = 896 instances)
  C5: synthetic 20-variable DAG, n=2000 rows subsampled to 1500 per run, nb_runs=16, 1+k candidates
Usage: python tools/bench_graph.py [train] [test] [exp_policy]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from copy import deepcopy
import numpy as np
import pandas as pd
from cgnn_b200 import Engine, synth, DirectedGraph
from cgnn_b200.plan import build_plan, n_params
from cgnn_b200.utils.scaling import scale

train = int(sys.argv[1]) if len(sys.argv) > 1 else 50
test = int(sys.argv[2]) if len(sys.argv) > 2 else 25
pol = int(sys.argv[3]) if len(sys.argv) > 3 else 3
EX = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "examples")
eng = Engine(0)


def candidates(dag):
    out = [dag]
    for a, b, _ in dag.get_list_edges():
        t = deepcopy(dag)
        t.reverse_edge(a, b)
        if not t.is_cyclic():
            out.append(t)
    return out


def flops_per_eval(d, train_step):      # fwd 3d+14, bwd +14+2d (non-symmetric PP: both as PT-like)
    return (5 * d + 28) if train_step else (3 * d + 14)


def run(name, data, graphs, n_runs, subsets=None, n_rows=None):
    cols = list(data.columns)
    plans = [build_plan(g, columns=cols) for g in graphs]
    X = scale(data.values).astype(np.float32)
    N = n_rows or X.shape[0]
    d = X.shape[1]
    kw = dict(n_runs=n_runs, h_layer_dim=20, train_epochs=train, test_epochs=test, exp_policy=pol, seed=1, row_subsets=subsets)
    eng.run_graph(X, plans[:2], **dict(kw, n_runs=2, row_subsets=None if subsets is None else subsets[:4]))     # warm-up
    r = eng.run_graph(X, plans, **kw)
    ms = eng.last_kernel_ms
    n_inst = len(plans) * n_runs
    evals = n_inst * (train + test) * 2.0 * N * N
    flops = n_inst * (train * 2.0 * N * N * flops_per_eval(d, True) + test * 2.0 * N * N * flops_per_eval(d, False))
    print("%s: %d graphs x %d runs = %d instances, N=%d d=%d, %d+%d steps: kernel %.1f ms -> %.0f train-steps/s, "
          "%.3g kernel-evals/s, %.2f algorithmic TFLOP/s, %.3g exp/s (7/eval), mean score %.4f, params %d" % (
              name, len(plans), n_runs, n_inst, N, d, train, test, ms, n_inst * train / (ms * 1e-3), evals / (ms * 1e-3),
              flops / (ms * 1e-3) / 1e12, 7 * evals / (ms * 1e-3), float(np.mean(r["scores"])), n_params(plans[0], 20)))


data = pd.read_csv(os.path.join(EX, "Example_graph_numdata.csv"))
dag = DirectedGraph(pd.read_csv(os.path.join(EX, "Example_graph_target.csv")))
run("C2", data, candidates(dag), 32)

X, edges = synth.make_dag_dataset(20, 2000, 3)
df = pd.DataFrame(X, columns=["V%d" % k for k in range(20)])
g = DirectedGraph()
for a, b in edges:
    g.add("V%d" % a, "V%d" % b, 1.0)
cands = candidates(g)[:10]
subs = np.stack([np.random.default_rng(k).permutation(2000)[:1500] for k in range(len(cands) * 16)]).astype(np.int32)
run("C5", df, cands, 16, subsets=subs, n_rows=1500)
print("peaks: mufu %.3g exp/s, ffma %.3g flop/s" % (eng.probe_peak("mufu"), eng.probe_peak("ffma")))
