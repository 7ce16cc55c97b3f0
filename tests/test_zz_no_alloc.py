"""The allocation-free entry point (cgnn_b200_run_pairwise_noalloc) against the ordinary one.

Collected last on purpose (file name): a run that stops at the first failure has then already run
every other test.  The check: for a caller that owns its GPU buffers and may not allocate - its
memory all sits in another framework's pool, or it runs under a latency budget - the library's
built-in scratch array must give the very same scores as cgnn_b200_run_pairwise, on the first call
in a process and on a following call that finds the previous call's leftovers in that array.
"""
import numpy as np
import pytest

from util import make_pair

pytestmark = pytest.mark.gpu
TOL = 1e-5      # the tolerance of tests/test_pairwise_gpu.py; the two paths are expected to agree exactly


def test_noalloc_matches_run_pairwise_twice(engine):
    import torch
    from cgnn_b200.noalloc import run_pairwise_noalloc
    from cgnn_b200 import EngineError

    calls = [  # (pairs, rows, runs, kwargs, uid_base): the second call is smaller and differently seeded,
               # so stale queue heads / table entries of the first would show
        (3, 200, 4, dict(h_layer_dim=12, train_epochs=25, test_epochs=6, seed=11, exp_policy=1), 0),
        (2, 333, 2, dict(h_layer_dim=20, train_epochs=15, test_epochs=5, seed=5, exp_policy=3), 96),
    ]
    for pairs, rows, runs, kw, uid in calls:
        xy = np.stack([make_pair(rows, 70 + p, p % 3) for p in range(pairs)])
        ref = engine.run_pairwise(xy, n_runs=runs, uid_base=uid, **kw)["scores"]
        d_xy = torch.from_numpy(xy).cuda()
        d_sc = torch.full((pairs, runs, 2), float("nan"), dtype=torch.float32, device="cuda")
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated()
        run_pairwise_noalloc(d_xy, d_sc, n_runs=runs, uid_base=uid, **kw)
        assert torch.cuda.memory_allocated() == before            # torch's allocator was not touched either
        got = d_sc.cpu().numpy()
        assert np.all(np.isfinite(got))
        np.testing.assert_allclose(got, ref, rtol=TOL, atol=0)
        assert np.array_equal(got, ref), "same kernel, same Philox streams: expected bit-identical scores"

    # beyond the stated maximum: a clear error, nothing launched
    big = torch.zeros((1, 64, 2), dtype=torch.float32, device="cuda")
    out = torch.zeros((1, 5000, 2), dtype=torch.float32, device="cuda")
    with pytest.raises(EngineError, match="exceeds the static scratch"):
        run_pairwise_noalloc(big, out, n_runs=5000, train_epochs=1, test_epochs=1)
    # host memory is refused (staging it would need an allocation)
    with pytest.raises(ValueError):
        run_pairwise_noalloc(torch.zeros((1, 64, 2)), out[:, :1], n_runs=1)


def test_noalloc_c_level_refusals(engine):
    """Straight through ctypes, past the Python wrapper's own checks: host pointers are refused with CGNN_ERR_ARG and a
    message (staging them would need the allocation this entry point exists to avoid), and the calling thread keeps
    its current device."""
    import ctypes as C
    import torch
    from cgnn_b200 import _lib
    from cgnn_b200.engine import make_hparams
    lib = _lib.load()
    hp = make_hparams(h_layer_dim=8, train_epochs=1, test_epochs=1)
    host_xy = np.zeros((1, 64, 2), np.float32)
    d_sc = torch.zeros((1, 1, 2), dtype=torch.float32, device="cuda")
    err = C.create_string_buffer(256)
    dev_before = torch.cuda.current_device()
    rc = lib.cgnn_b200_run_pairwise_noalloc(0, C.c_void_p(0), C.byref(hp), C.c_void_p(host_xy.ctypes.data), 1, 64, 1,
                                            C.c_ulonglong(0), C.c_void_p(d_sc.data_ptr()), err, len(err))
    assert rc == -1 and b"device memory" in err.value                     # CGNN_ERR_ARG
    rc = lib.cgnn_b200_run_pairwise_noalloc(0, C.c_void_p(0), C.byref(hp), None, 1, 64, 1, C.c_ulonglong(0),
                                            C.c_void_p(d_sc.data_ptr()), err, len(err))
    assert rc == -1 and b"NULL" in err.value
    assert torch.cuda.current_device() == dev_before
