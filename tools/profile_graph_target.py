"""Small fixed workload for ncu: one launch of the graph instance kernel on the bundled 22-variable
example (148 instances = one wave), or with shape=c5 on the synthetic 20-variable, 1500-row shape of
BASELINE configs[4].  Writes gpurun_out/prof_meta_graph_<shape>.json for tools/ncu_summary.py --meta.
Usage: python tools/profile_graph_target.py [pol] [train] [test] [c2|c5]"""
import json
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, pandas as pd
from cgnn_b200 import Engine, DirectedGraph
from cgnn_b200.plan import build_plan
from cgnn_b200.utils.scaling import scale
pol = int(sys.argv[1]) if len(sys.argv) > 1 else 3
train = int(sys.argv[2]) if len(sys.argv) > 2 else 8
test = int(sys.argv[3]) if len(sys.argv) > 3 else 4
shape = sys.argv[4] if len(sys.argv) > 4 else "c2"
EX = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "examples")
data = pd.read_csv(os.path.join(EX, "Example_graph_numdata.csv"))
dag = DirectedGraph(pd.read_csv(os.path.join(EX, "Example_graph_target.csv")))
plan = build_plan(dag, columns=list(data.columns))
if shape == "c5":
    from cgnn_b200 import synth
    Xs, edges = synth.make_dag_dataset(20, 1500, 3)
    data = pd.DataFrame(Xs, columns=["V%d" % k for k in range(20)])
    dag = DirectedGraph()
    for a, b in edges:
        dag.add("V%d" % a, "V%d" % b, 1.0)
    plan = build_plan(dag, columns=list(data.columns))
X = scale(data.values).astype(np.float32)
eng = Engine(0)
for rep in range(2):
    r = eng.run_graph(X, [plan] * 4, n_runs=37, h_layer_dim=20, train_epochs=train, test_epochs=test, exp_policy=pol, seed=1)
    print("graph pol %d: 148 instances x (%d+%d) steps: kernel %.3f ms, mean score %.4f" % (pol, train, test, eng.last_kernel_ms, float(np.mean(r["scores"]))))
from cgnn_b200.build import source_hash
n, d = X.shape
# kernel evaluations the launch performs: gradient pass N^2 observed + N^2 generated columns per row sweep; loss pass
# N^2 + N(N+1)/2 (generated block taken once per unordered tile pair); the observed-observed block once per instance
evals = 148 * (train * 2 * n * n + test * (n * n + n * (n + 1) // 2) + n * (n + 1) // 2)
flops = 148 * (train * (n * n * (5 * d + 28) + (n * (n - 1) // 2) * (7 * d + 28)) + test * (3 * d + 14) * (n * n + n * (n - 1) // 2))
meta = dict(family="graph", shape=shape, exp_policy=pol, instances=148, rows=n, vars=d, h=20, train=train, test=test,
            evals_per_launch=evals, algorithmic_flops_per_launch=flops, source_hash=source_hash("graph"))
os.makedirs("gpurun_out", exist_ok=True)
json.dump(meta, open("gpurun_out/prof_meta_graph_%s.json" % shape, "w"))
