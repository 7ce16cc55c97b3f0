"""CPU stand-in for the reference's TensorFlow-CPU path (TEST / BENCH INFRASTRUCTURE ONLY).

TensorFlow is not installed on this image (nor on the GPU box) and cannot be installed, so the
reference's own CPU path cannot be timed.  What is timed instead, and labelled kind="port", is the
op-for-op fp32 torch-CPU restatement of the reference graph (oracle.gnn / oracle.mmd: the full
2N x 2N Gram expansion of Loss.py:12-32, seven exps, autograd, TF-style Adam), driven like the
reference: one Python-level call per step (GNN.py:92-96) and one process per run under
joblib.Parallel (GNN.py:193-194) with one torch thread per worker.
"""
import os
import time

import numpy as np


def _one_instance(data, h, train_steps, eval_steps, seed):
    import torch
    torch.set_num_threads(1)
    from . import gnn
    rng = np.random.default_rng(seed)
    n = data.shape[0]
    theta0 = (0.05 * rng.normal(size=gnn.n_params(h))).astype(np.float32)
    ntr = rng.normal(size=(train_steps, n)).astype(np.float32)
    nev = rng.normal(size=(eval_steps, n)).astype(np.float32)
    t0 = time.perf_counter()
    theta, _, _ = gnn.train(theta0, data.astype(np.float32), ntr, h, dtype=torch.float32)
    t1 = time.perf_counter()
    score, _ = gnn.evaluate(theta, data.astype(np.float32), nev, h, dtype=torch.float32)
    t2 = time.perf_counter()
    return t1 - t0, t2 - t1, score


def time_pairwise_sample(data, h=30, n_instances=None, train_steps=8, eval_steps=4, n_jobs=None):
    """Runs n_instances independent (train_steps + eval_steps)-step instances over n_jobs processes.
    Returns dict(train_steps_per_s, wall_s, cores, sample)."""
    from joblib import Parallel, delayed
    n_jobs = n_jobs or os.cpu_count() or 1
    n_instances = n_instances or n_jobs
    t0 = time.perf_counter()
    res = Parallel(n_jobs=n_jobs)(delayed(_one_instance)(data, h, train_steps, eval_steps, 1000 + k)
                                  for k in range(n_instances))
    wall = time.perf_counter() - t0
    ttr = sum(r[0] for r in res)
    tev = sum(r[1] for r in res)
    # throughput in the benchmark's unit: train steps per second of a job whose instances carry the
    # same train:eval mix as the real workload (1000:500) - scale the measured per-step costs.
    per_train = ttr / (n_instances * train_steps)
    per_eval = tev / max(1, n_instances * eval_steps)
    return dict(per_train_step_s=per_train, per_eval_step_s=per_eval, wall_s=wall, cores=n_jobs,
                n_instances=n_instances, train_steps=train_steps, eval_steps=eval_steps)
