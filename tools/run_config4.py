"""BASELINE.json configs[3] in full: the synthetic pairwise sweep - 1000 cause-effect pairs, n=1500 rows,
nb_runs=32, h=30, 1000 train + 500 eval steps, both directions (64,000 generator instances, 64 M train steps) -
through the public API (cgnn_b200.GNN.predict_dataset) on all GPUs of a torchrun job.
Usage: torchrun --nproc-per-node N tools/run_config4.py [n_pairs] [out.json]"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import pandas as pd
import torch, torch.distributed as dist
from sklearn.metrics import average_precision_score, roc_auc_score
import cgnn_b200 as cgnn
from cgnn_b200 import synth

n_pairs = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
out = sys.argv[2] if len(sys.argv) > 2 else "gpurun_out/r2_config4.json"
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
S = cgnn.SETTINGS
S.verbose = False
S.h_layer_dim, S.NB_RUNS, S.train_epochs, S.test_epochs = 30, 32, 1000, 500
xy, labels = synth.make_pairs(0, n_pairs, 1500)
df = pd.DataFrame({"SampleID": ["pair%d" % k for k in range(n_pairs)],
                   "A": [xy[k, :, 0].astype(np.float64) for k in range(n_pairs)],
                   "B": [xy[k, :, 1].astype(np.float64) for k in range(n_pairs)]})
model = cgnn.GNN(backend="B200")
model.predict_dataset(df.iloc[:world])                 # warm-up: context, module load
if world > 1:
    dist.barrier()
torch.cuda.synchronize()
t0 = time.perf_counter()
pred = np.array(model.predict_dataset(df))
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
dt = time.perf_counter() - t0
if rank == 0:
    y = (labels > 0).astype(int)
    res = dict(config="BASELINE configs[3]: synthetic pairwise sweep", n_pairs=n_pairs, n_gpus=world, rows=1500, nb_runs=32, h=30,
               train_epochs=1000, test_epochs=500, instances=n_pairs * 64, seconds=dt, pair_scores_per_s=n_pairs / dt,
               train_steps_per_s=n_pairs * 64 * 1000 / dt, accuracy=float(np.mean(np.sign(pred) == labels)),
               aupr=float(average_precision_score(y, pred)), auroc=float(roc_auc_score(y, pred)),
               api="cgnn_b200.GNN.predict_dataset (host DataFrame in, predictions out)")
    print(json.dumps(res))
    json.dump(res, open(out, "w"), indent=1)
if world > 1:
    dist.destroy_process_group()
