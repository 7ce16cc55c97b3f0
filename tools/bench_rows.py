"""Per-evaluation speed of the pairwise instance kernel against the pair length (short pairs are what real cause-effect
collections are made of): one launch of enough instances to fill the GPU several times over for every row count, with
the build chosen by the library's per-pair rule or forced through CGNN_PW_SPLIT (set before the first call).
Usage: python tools/bench_rows.py [train] [test]        e.g.  CGNN_PW_SPLIT=0,0 python tools/bench_rows.py   (all on 512 threads)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from cgnn_b200 import Engine, synth

train = int(sys.argv[1]) if len(sys.argv) > 1 else 200
test = int(sys.argv[2]) if len(sys.argv) > 2 else 100
eng = Engine(0)
base = None
print("CGNN_PW_SPLIT =", os.environ.get("CGNN_PW_SPLIT", "(library default)"))
for n in (1500, 1000, 700, 640, 512, 400, 300, 250, 200, 160, 128, 100, 64, 48):
    per_inst = n * n * 1.5
    n_pairs = int(min(600, max(10, round(9 * 1500 * 1500 * 1.5 / per_inst))))     # about the work of 9 full-size pairs
    xy = np.stack([synth.make_pair(k, n)[0] for k in range(n_pairs)])
    kw = dict(n_runs=32, h_layer_dim=30, train_epochs=train, test_epochs=test, seed=1)
    eng.run_pairwise(xy[:2], **dict(kw, n_runs=2))
    eng.run_pairwise(xy, **kw)
    ms = eng.last_kernel_ms
    evals = n_pairs * 64 * (train + test) * (n * n + n * (n - 1) / 2.0)
    rate = evals / (ms * 1e-3)
    base = base or rate
    print("rows %5d: %4d pairs x 64 instances, %8.2f ms, %.3e kernel evaluations/s = %.2f of the 1500-row rate, %.0f train-steps/s"
          % (n, n_pairs, ms, rate, rate / base, n_pairs * 64 * train / (ms * 1e-3)))
