"""Where instances run: GPUs of this process and ranks of a torch.distributed job.

The reference spreads the `nb_runs` restarts of ONE pair / candidate over a joblib process pool and
pins run r to `/gpu:(gpu_offset + r % nb_gpu)` (GNN.py:158-163,193-194; CGNN.py:187-191,214-215).
Here a whole batch of independent instances is cut into contiguous shards:

  * over the ranks of an initialised torch.distributed job (one process per GPU, `torchrun`), then
  * over the GPUs this process drives (SETTINGS.GPU_OFFSET .. +SETTINGS.NB_GPU, one Engine each, one
    host thread per Engine - the C call releases the GIL),

and the score arrays are put back together with ONE all_gather (NCCL on GPUs, gloo in the CPU
tests).  No other collective exists on this path.  Every instance draws its random numbers from the
Philox stream of its GLOBAL index (`uid_base`), so the result is bit-identical for any sharding.

Two schedules:
  * static (default): equal contiguous shards - optimal when all instances cost the same (one launch
    per GPU, every SM busy until the common end);
  * work stealing (`SETTINGS.work_stealing`, or automatically for batches whose items differ in cost):
    the batch is cut into chunks; every worker (rank x local GPU) starts on "its" chunk - the static
    round-robin part - and then pulls the next unclaimed chunk from a shared counter: an atomic `add` on
    the job's c10d store across processes, a locked integer across the GPUs of one process.  Results are
    merged with one all_reduce(SUM) over a zero-initialised buffer (each item is written by one worker).
"""
import itertools
import os
import threading
from concurrent.futures import ThreadPoolExecutor

import numpy as np

from .utils.Settings import SETTINGS

_ENGINES = {}
_LOCK = threading.Lock()


def split_range(n, parts, k):
    """[lo, hi) of the k-th of `parts` near-equal contiguous shards of range(n)."""
    base, extra = divmod(n, parts)
    lo = k * base + min(k, extra)
    return lo, lo + base + (1 if k < extra else 0)


def dist_info():
    """(rank, world) of the torch.distributed job, (0, 1) when there is none."""
    try:
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            return dist.get_rank(), dist.get_world_size()
    except ImportError:
        pass
    return 0, 1


def local_devices():
    """CUDA ordinals this process drives.  Under torchrun: LOCAL_RANK only."""
    _, world = dist_info()
    if world > 1:
        return [int(os.environ.get("LOCAL_RANK", "0"))]
    return [int(SETTINGS.GPU_OFFSET) + k for k in range(max(1, int(SETTINGS.NB_GPU)))]


def get_engines():
    from .engine import Engine
    out = []
    with _LOCK:
        for d in local_devices():
            if d not in _ENGINES:
                _ENGINES[d] = Engine(d)
            out.append(_ENGINES[d])
    return out


def close_engines():
    with _LOCK:
        for e in _ENGINES.values():
            e.close()
        _ENGINES.clear()


def all_gather_concat(local, counts):
    """Concatenate per-rank float32 arrays (rank k contributes counts[k] leading rows) on every rank."""
    rank, world = dist_info()
    if world == 1:
        return local
    import torch
    import torch.distributed as dist
    backend = dist.get_backend()
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0"))) if backend == "nccl" else torch.device("cpu")
    tail = tuple(local.shape[1:])
    width = max(counts)
    buf = torch.zeros((width,) + tail, dtype=torch.float32, device=dev)
    buf[:local.shape[0]] = torch.from_numpy(np.ascontiguousarray(local, np.float32)).to(dev)
    parts = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(parts, buf)                                   # the path's only collective
    return np.concatenate([p[:c].cpu().numpy() for p, c in zip(parts, counts)], axis=0)


def run_sharded(n_items, items_per_uid, uid_base, call, tail_shape=()):
    """Cut range(n_items) over ranks then local GPUs; `call(engine, lo, hi, uid)` must return a float32
    array whose first axis has hi-lo rows.  Returns the full [n_items, ...] array on every rank."""
    rank, world = dist_info()
    counts = [split_range(n_items, world, k)[1] - split_range(n_items, world, k)[0] for k in range(world)]
    r_lo, r_hi = split_range(n_items, world, rank)
    engines = get_engines()
    jobs = []
    for k, eng in enumerate(engines):
        lo, hi = split_range(r_hi - r_lo, len(engines), k)
        if hi > lo:
            jobs.append((eng, r_lo + lo, r_lo + hi))
    if not jobs:
        local = None
    elif len(jobs) == 1:
        e, lo, hi = jobs[0]
        local = call(e, lo, hi, uid_base + lo * items_per_uid)
    else:
        with ThreadPoolExecutor(len(jobs)) as ex:
            parts = list(ex.map(lambda j: call(j[0], j[1], j[2], uid_base + j[1] * items_per_uid), jobs))
        local = np.concatenate(parts, axis=0)
    if world == 1:
        return local
    if local is None:                                          # more ranks than items
        local = np.zeros((0,) + tuple(tail_shape), np.float32)
    return all_gather_concat(local, counts)


_CALL_SEQ = itertools.count()


class _ChunkQueue(object):
    """Next-unclaimed-chunk counter shared by every worker of the job."""

    def __init__(self, n_chunks, n_workers_total, first_worker, n_local):
        self.n_chunks, self.total = n_chunks, n_workers_total
        self.first, self.n_local = first_worker, n_local
        self.key = "cgnn_b200/steal/%d" % next(_CALL_SEQ)         # all ranks call in lockstep -> same key
        self.lock = threading.Lock()
        self.local_next = n_workers_total                          # chunks 0..total-1 are the static first picks
        _, world = dist_info()
        self.store = None
        if world > 1:
            import torch.distributed as dist
            self.store = dist.distributed_c10d._get_default_store()

    def first_pick(self, local_worker):
        c = self.first + local_worker
        return c if c < self.n_chunks else None

    def steal(self):
        if self.store is not None:
            c = self.total + self.store.add(self.key, 1) - 1       # atomic across processes
        else:
            with self.lock:
                c = self.local_next
                self.local_next += 1
        return c if c < self.n_chunks else None


def run_stolen(n_items, items_per_uid, uid_base, call, tail_shape=(), chunk_items=None):
    """Like run_sharded, with the chunked work-stealing schedule.  `chunk_items`: items per chunk (default:
    about four chunks per worker)."""
    rank, world = dist_info()
    engines = get_engines()
    n_local = len(engines)
    total = world * n_local
    if chunk_items is None:
        chunk_items = max(1, -(-n_items // (4 * total)))
    n_chunks = -(-n_items // chunk_items)
    queue = _ChunkQueue(n_chunks, total, rank * n_local, n_local)
    out = np.zeros((n_items,) + tuple(tail_shape), np.float32)
    owned = []

    def worker(k):
        c = queue.first_pick(k)
        while c is not None:
            lo, hi = c * chunk_items, min(n_items, (c + 1) * chunk_items)
            out[lo:hi] = call(engines[k], lo, hi, uid_base + lo * items_per_uid)
            owned.append(c)
            c = queue.steal()

    if n_local == 1:
        worker(0)
    else:
        with ThreadPoolExecutor(n_local) as ex:
            list(ex.map(worker, range(n_local)))
    run_stolen.last_owned = sorted(owned)
    if world == 1:
        return out
    import torch
    import torch.distributed as dist
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0"))) if dist.get_backend() == "nccl" else torch.device("cpu")
    buf = torch.from_numpy(out).to(dev)
    dist.all_reduce(buf, op=dist.ReduceOp.SUM)                      # every item was written by exactly one worker
    return buf.cpu().numpy()


def run_batch(n_items, items_per_uid, uid_base, call, tail_shape=(), costs=None):
    """Static shards for uniform batches, work stealing when asked for or when item costs differ."""
    steal = bool(getattr(SETTINGS, "work_stealing", False))
    if costs is not None and len(costs) > 1 and max(costs) > 1.25 * min(costs):
        steal = True
    _, world = dist_info()
    if steal and (world > 1 or len(local_devices()) > 1):
        return run_stolen(n_items, items_per_uid, uid_base, call, tail_shape)
    return run_sharded(n_items, items_per_uid, uid_base, call, tail_shape)
