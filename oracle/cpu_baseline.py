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


def _one_instance(data, h, train_steps, eval_steps, seed, repeats=1, warmup_steps=1):
    """One worker process: `warmup_steps` untimed train steps (first-call costs of torch / autograd), then
    `repeats` timed blocks of train_steps + eval_steps.  Returns [(train_s, eval_s)] per block and the last score."""
    import torch
    torch.set_num_threads(1)
    from . import gnn
    rng = np.random.default_rng(seed)
    n = data.shape[0]
    d32 = data.astype(np.float32)
    theta = (0.05 * rng.normal(size=gnn.n_params(h))).astype(np.float32)
    if warmup_steps > 0:
        gnn.train(theta, d32, rng.normal(size=(warmup_steps, n)).astype(np.float32), h, dtype=torch.float32)
    blocks, score = [], float("nan")
    for _ in range(repeats):
        ntr = rng.normal(size=(train_steps, n)).astype(np.float32)
        nev = rng.normal(size=(eval_steps, n)).astype(np.float32)
        t0 = time.perf_counter()
        theta, _, _ = gnn.train(theta, d32, ntr, h, dtype=torch.float32)
        t1 = time.perf_counter()
        score, _ = gnn.evaluate(theta, d32, nev, h, dtype=torch.float32)
        t2 = time.perf_counter()
        blocks.append((t1 - t0, t2 - t1))
    return blocks, score


def time_pairwise_sample(data, h=30, n_instances=None, train_steps=3, eval_steps=2, n_jobs=None, repeats=3,
                         warmup_steps=1):
    """n_instances independent instances over n_jobs processes, each timing `repeats` blocks of
    (train_steps + eval_steps) steps after an untimed warm-up.  Per block the per-step costs are averaged over the
    workers; the caller turns them into a throughput and takes the median over the blocks.
    Returns dict(blocks=[(per_train_step_s, per_eval_step_s)], per_train_step_s, per_eval_step_s (medians), ...)."""
    from joblib import Parallel, delayed
    n_jobs = n_jobs or os.cpu_count() or 1
    n_instances = n_instances or n_jobs
    t0 = time.perf_counter()
    res = Parallel(n_jobs=n_jobs)(delayed(_one_instance)(data, h, train_steps, eval_steps, 1000 + k, repeats, warmup_steps)
                                  for k in range(n_instances))
    wall = time.perf_counter() - t0
    blocks = []
    for r in range(repeats):
        ttr = sum(b[0][r][0] for b in res)
        tev = sum(b[0][r][1] for b in res)
        blocks.append((ttr / (n_instances * train_steps), tev / max(1, n_instances * eval_steps)))
    return dict(blocks=blocks, per_train_step_s=float(np.median([b[0] for b in blocks])),
                per_eval_step_s=float(np.median([b[1] for b in blocks])), wall_s=wall, cores=n_jobs,
                n_instances=n_instances, train_steps=train_steps, eval_steps=eval_steps, repeats=repeats,
                warmup_steps=warmup_steps)
