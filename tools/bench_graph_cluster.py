"""Per-instance time of the graph kernel against the cluster size, for batches too small to fill the GPU with one
CTA per instance (hill-climbing tails; candidate lists sharded over 8 GPUs).  Bundled 22-variable shape (N=500) and
the 20-variable x 1500-row shape.  Usage: python tools/bench_graph_cluster.py [n_instances] [train] [test]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import pandas as pd
from cgnn_b200 import Engine, synth, DirectedGraph
from cgnn_b200.plan import build_plan
from cgnn_b200.utils.scaling import scale

n_inst = int(sys.argv[1]) if len(sys.argv) > 1 else 36
train = int(sys.argv[2]) if len(sys.argv) > 2 else 40
test = int(sys.argv[3]) if len(sys.argv) > 3 else 20
EX = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "examples")
eng = Engine(0)


def run(name, X, plan):
    base = None
    for c in (0, 1, 2, 3, 4, 5, 6, 8):
        kw = dict(n_runs=n_inst, h_layer_dim=20, train_epochs=train, test_epochs=test, seed=1, graph_cluster=c)
        eng.run_graph(X, [plan], **kw)
        r = eng.run_graph(X, [plan], **kw)
        ms = eng.last_kernel_ms
        if c == 1:
            base = (ms, r["scores"].copy())
        same = "" if base is None or c < 2 else ("  scores equal to 1 CTA: %s" % np.array_equal(base[1], r["scores"]))
        print("%s: %d instances, graph_cluster=%d (used %d): kernel %.2f ms%s%s" % (
            name, n_inst, c, eng.last_graph_cluster, ms, "" if base is None or c < 2 else "  x%.2f vs 1 CTA" % (base[0] / ms), same))


data = pd.read_csv(os.path.join(EX, "Example_graph_numdata.csv"))
dag = DirectedGraph(pd.read_csv(os.path.join(EX, "Example_graph_target.csv")))
run("C2 shape (N=500, 22 vars)", scale(data.values).astype(np.float32), build_plan(dag, columns=list(data.columns)))
Xs, edges = synth.make_dag_dataset(20, 1500, 3)
df = pd.DataFrame(Xs, columns=["V%d" % k for k in range(20)])
g = DirectedGraph()
for a, b in edges:
    g.add("V%d" % a, "V%d" % b, 1.0)
run("C5 shape (N=1500, 20 vars)", scale(df.values).astype(np.float32), build_plan(g, columns=list(df.columns)))
