"""Small, fixed workload for ncu: one launch of the pairwise instance kernel (one wave of instances,
a few steps each).  Writes gpurun_out/prof_meta_pairwise.json (evaluations per launch, source hash) for
tools/ncu_summary.py --meta.  Usage: python tools/profile_target.py [exp_policy] [train] [test] [pairs] [n]"""
import json
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from cgnn_b200 import Engine, synth

pol = int(sys.argv[1]) if len(sys.argv) > 1 else 1
train = int(sys.argv[2]) if len(sys.argv) > 2 else 20
test = int(sys.argv[3]) if len(sys.argv) > 3 else 10
pairs = int(sys.argv[4]) if len(sys.argv) > 4 else 37      # 37 * 8 = 296 instances = one full wave
n = int(sys.argv[5]) if len(sys.argv) > 5 else 1500
xy, _ = synth.make_pairs(0, pairs, n)
eng = Engine(0)
for rep in range(2):
    r = eng.run_pairwise(xy, n_runs=4, h_layer_dim=30, train_epochs=train, test_epochs=test, exp_policy=pol, seed=1)
    print("pol %d: %d instances x (%d+%d) steps: kernel %.3f ms, mean score %.5f" % (
        pol, pairs * 8, train, test, eng.last_kernel_ms, float(np.mean(r["scores"]))))
from cgnn_b200.build import source_hash
n_inst = pairs * 8
P = n * n + n * (n - 1) // 2                  # kernel evaluations of one step, + the once-per-instance TT block
meta = dict(family="pairwise", exp_policy=pol, instances=n_inst, rows=n, h=30, train=train, test=test,
            evals_per_launch=n_inst * (P * (train + test) + n * (n + 1) // 2), source_hash=source_hash("pairwise"))
os.makedirs("gpurun_out", exist_ok=True)
json.dump(meta, open("gpurun_out/prof_meta_pairwise.json", "w"))
