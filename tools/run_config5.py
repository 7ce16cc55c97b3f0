"""BASELINE.json configs[4]: synthetic 20-variable random DAG, n=2000 rows (each run subsamples 1500,
CGNN.py:183-185), nb_runs=16, h=20: orient the true skeleton with the pairwise GNN, then CGNN hill climbing
with the candidate graphs sharded over the GPUs of a torchrun job.
Usage: torchrun --nproc-per-node N tools/run_config5.py [out.json]"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import pandas as pd
import torch, torch.distributed as dist
import cgnn_b200 as cgnn
from cgnn_b200 import synth
from cgnn_b200.CGNN import hill_climbing

out = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/r2_config5.json"
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
S = cgnn.SETTINGS
S.verbose = False
S.NB_RUNS, S.h_layer_dim, S.train_epochs, S.test_epochs = 16, 20, 1000, 500
X, edges = synth.make_dag_dataset(20, 2000, 3)
df = pd.DataFrame(X, columns=["V%d" % k for k in range(20)])
truth = {("V%d" % a, "V%d" % b) for a, b in edges}
umg = cgnn.UndirectedGraph(pd.DataFrame(sorted(truth), columns=["Node1", "Node2"]))
t0 = time.perf_counter()
dag = cgnn.GNN(backend="B200").orient_graph(df, umg)
t1 = time.perf_counter()
pw_hits = sum((a, b) in truth for a, b, _ in dag.get_list_edges())
res_graph = cgnn.CGNN(backend="B200").orient_directed_graph(df, dag)
t2 = time.perf_counter()
if rank == 0:
    hits = sum((a, b) in truth for a, b, _ in res_graph.get_list_edges())
    n_graphs = int(sum(hill_climbing.last_batches))
    res = dict(config="BASELINE configs[4]: synthetic 20-variable DAG, n=2000 -> 1500 per run, nb_runs=16", n_gpus=world,
               n_edges=len(truth), pairwise_correct=int(pw_hits), cgnn_correct=int(hits), acyclic=not res_graph.is_cyclic(),
               seconds_pairwise=t1 - t0, seconds_hill_climbing=t2 - t1, batches=list(hill_climbing.last_batches),
               candidate_graphs_scored=n_graphs, candidate_graphs_per_s=n_graphs / (t2 - t1),
               graph_train_steps_per_s=n_graphs * 16 * 1000 / (t2 - t1))
    print(json.dumps(res))
    json.dump(res, open(out, "w"), indent=1)
if world > 1:
    dist.destroy_process_group()
