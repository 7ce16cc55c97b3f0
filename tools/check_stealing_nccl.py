"""torchrun check of the chunked work-stealing schedule on real GPUs (NCCL): ragged pairs are scored by N ranks
pulling chunks from the c10d store; every rank must end with the scores of a single-GPU static run."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import torch, torch.distributed as dist
from util import make_pair
from cgnn_b200 import Engine, SETTINGS, runtime
from cgnn_b200.GNN import score_pairs

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
SETTINGS.verbose = False
pairs = [make_pair(64 + 48 * (k % 6), 900 + k, k % 3) for k in range(40)]
kw = dict(nb_runs=2, h_layer_dim=8, train_epochs=30, test_epochs=10, seed=5)
got = score_pairs(pairs, 0, **kw)                       # ragged -> stealing schedule over the ranks
owned = list(runtime.run_stolen.last_owned)
# reference: this rank alone, static
eng = Engine(local)
ld = max(p.shape[0] for p in pairs)
xy = np.zeros((len(pairs), ld, 2), np.float32)
for k, p in enumerate(pairs):
    xy[k, :p.shape[0]] = p
want = eng.run_pairwise(xy, n_rows=np.array([p.shape[0] for p in pairs], np.int32), n_runs=2, h_layer_dim=8,
                        train_epochs=30, test_epochs=10, seed=5)["scores"]
ok = np.array_equal(got, want)
gathered = [None] * world
dist.all_gather_object(gathered, (rank, owned, bool(ok)))
if rank == 0:
    print(gathered)
    chunks = sorted(c for _, o, _ in gathered for c in o)
    assert all(g[2] for g in gathered), "scores differ from the single-GPU run"
    assert chunks == list(range(len(chunks))) and len(chunks) >= 2 * world
    print("work stealing over %d GPUs: OK (%d chunks)" % (world, len(chunks)))
dist.destroy_process_group()
