"""N>1 host path on CPU: two gloo ranks shard a batch, score their shards on fake engines and
all_gather - every rank must end with the single-process result."""
import os
import socket
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    sys.path.insert(0, HERE)
    sys.path.insert(0, os.path.dirname(HERE))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import pandas as pd
    from fake_engine import FakeEngine
    from cgnn_b200 import runtime, SETTINGS
    from cgnn_b200.GNN import score_pairs
    from cgnn_b200.search import GraphScorer
    import test_host_search as ths
    import cgnn_b200.utils.Graph as my_graph
    SETTINGS.verbose = False
    eng = FakeEngine(rank)
    runtime.get_engines = lambda: [eng]
    rng = np.random.default_rng(0)
    pairs = [rng.normal(size=(60, 2)).astype(np.float32) for k in range(5)]          # uniform cost -> static shards
    s = score_pairs(pairs, 0, nb_runs=3)
    g = ths.build_start(my_graph, False)
    from cgnn_b200.CGNN import hill_climbing
    out = hill_climbing(g, ths._data(False), nb_runs=2)
    q.put((rank, s, ths._edges(out), [c for c in eng.calls if c[0] == "pairwise"]))
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_two_gloo_ranks_reproduce_single_process(monkeypatch):
    import torch.multiprocessing as mp
    sys.path.insert(0, HERE)
    from fake_engine import FakeEngine
    from cgnn_b200 import runtime, SETTINGS
    from cgnn_b200.GNN import score_pairs
    import test_host_search as ths
    import cgnn_b200.utils.Graph as my_graph
    from cgnn_b200.CGNN import hill_climbing
    monkeypatch.setattr(SETTINGS, "verbose", False)
    monkeypatch.setattr(runtime, "get_engines", lambda: [FakeEngine()])
    rng = np.random.default_rng(0)
    pairs = [rng.normal(size=(60, 2)).astype(np.float32) for k in range(5)]
    want = score_pairs(pairs, 0, nb_runs=3)
    want_edges = ths._edges(hill_climbing(ths.build_start(my_graph, False), ths._data(False), nb_runs=2))

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    for rank, s, edges, calls in res:
        assert np.array_equal(s, want), "rank %d: sharded scores differ from the single-process scores" % rank
        assert edges == want_edges
    by_rank = {r: calls for r, _, _, calls in res}
    assert by_rank[0][0][1] == 3 and by_rank[1][0][1] == 2          # 5 pairs -> shards of 3 and 2
    assert by_rank[1][0][3] == 3 * 3 * 2                           # rank 1 starts at its global instance index


def test_split_range():
    from cgnn_b200.runtime import split_range
    for n in (0, 1, 5, 16, 17):
        for parts in (1, 2, 3, 8):
            cuts = [split_range(n, parts, k) for k in range(parts)]
            assert cuts[0][0] == 0 and cuts[-1][1] == n
            assert all(cuts[k][1] == cuts[k + 1][0] for k in range(parts - 1))
            sizes = [b - a for a, b in cuts]
            assert max(sizes) - min(sizes) <= 1


def _steal_worker(rank, world, port, q):
    sys.path.insert(0, HERE)
    sys.path.insert(0, os.path.dirname(HERE))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    import time
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from fake_engine import FakeEngine
    from cgnn_b200 import runtime, SETTINGS
    from cgnn_b200.GNN import score_pairs
    SETTINGS.verbose = False

    class Slow(FakeEngine):
        def run_pairwise(self, *a, **k):
            time.sleep(0.25 if rank == 0 else 0.0)          # rank 0 is the slow GPU: rank 1 must take more chunks
            return FakeEngine.run_pairwise(self, *a, **k)

    eng = Slow(rank)
    runtime.get_engines = lambda: [eng]
    rng = np.random.default_rng(0)
    pairs = [rng.normal(size=(40 + 30 * (k % 5), 2)).astype(np.float32) for k in range(24)]      # ragged -> stealing schedule
    s = score_pairs(pairs, 0, nb_runs=2)
    q.put((rank, s, list(runtime.run_stolen.last_owned)))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_work_stealing_over_two_ranks(monkeypatch):
    """Ragged batch -> chunked schedule: static first pick + atomic counter on the c10d store.  Every chunk is
    done exactly once, the fast rank does more of them, and the merged scores equal the single-process ones."""
    import torch.multiprocessing as mp
    sys.path.insert(0, HERE)
    from fake_engine import FakeEngine
    from cgnn_b200 import runtime, SETTINGS
    from cgnn_b200.GNN import score_pairs
    monkeypatch.setattr(SETTINGS, "verbose", False)
    monkeypatch.setattr(runtime, "get_engines", lambda: [FakeEngine()])
    rng = np.random.default_rng(0)
    pairs = [rng.normal(size=(40 + 30 * (k % 5), 2)).astype(np.float32) for k in range(24)]
    want = score_pairs(pairs, 0, nb_runs=2)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_steal_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    owned = {r: o for r, _, o in res}
    assert sorted(owned[0] + owned[1]) == list(range(8)), owned            # 24 items / (4 chunks x 2 workers) = 3 per chunk
    assert owned[0][0] == 0 and owned[1][0] == 1                           # static first picks
    assert len(owned[1]) > len(owned[0])                                   # the fast rank stole more
    for _, s, _ in res:
        assert np.array_equal(s, want)
