"""Python face of cgnn_b200_run_pairwise_noalloc (include/cgnn_b200.h): score a batch of pairs
without a single allocation inside the library.

Meant for hosts that cannot call cudaMalloc while they work - all device memory already belongs
to another framework's pool, the calling loop has a latency budget that a device-synchronising
allocation would break, or the process froze its memory map after start-up.  You bring the input
and output tensors (already on the GPU); the library brings a fixed scratch array that is part of
its image.  One call at a time per process: the scratch is shared.
"""
import ctypes as C

from . import _lib
from .engine import EngineError, make_hparams
from .utils.Settings import SETTINGS


def run_pairwise_noalloc(xy, scores, n_runs=None, uid_base=0, device=None, stream=None, **kw):
    """xy: cuda float32 tensor [n_pairs, n_rows, 2]; scores: cuda float32 tensor [n_pairs, n_runs, 2]
    that receives the result.  Neither tensor is copied, resized or re-allocated.  Same scores, bit
    for bit, as Engine.run_pairwise on the same seed / uid_base."""
    lib = _lib.load()
    hp = make_hparams(**kw)
    n_runs = int(n_runs if n_runs is not None else kw.get("nb_runs", SETTINGS.NB_RUNS))
    if not (xy.is_cuda and scores.is_cuda and xy.is_contiguous() and scores.is_contiguous()):
        raise ValueError("xy and scores must be contiguous CUDA tensors")
    n_pairs, n_rows = int(xy.shape[0]), int(xy.shape[1])
    if tuple(scores.shape) != (n_pairs, n_runs, 2):
        raise ValueError("scores must be [n_pairs, n_runs, 2]")
    dev = xy.device.index if device is None else int(device)
    err = C.create_string_buffer(512)
    rc = lib.cgnn_b200_run_pairwise_noalloc(dev, C.c_void_p(stream or 0), C.byref(hp), C.c_void_p(xy.data_ptr()),
                                            n_pairs, n_rows, n_runs, C.c_ulonglong(int(uid_base)),
                                            C.c_void_p(scores.data_ptr()), err, len(err))
    if rc != 0:
        raise EngineError("cgnn_b200_run_pairwise_noalloc failed (%d): %s" % (rc, err.value.decode()))
    return scores
