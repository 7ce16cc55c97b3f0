"""ctypes binding of libcgnn_b200.so (include/cgnn_b200.h).  No fallback: if the library is
missing or no sm_100 GPU is present the product raises."""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("CGNN_B200_LIB") or os.path.join(HERE, "libcgnn_b200.so")      # env override: tuning experiments


class Hparams(C.Structure):
    _fields_ = [("h_layer_dim", C.c_int), ("train_epochs", C.c_int), ("test_epochs", C.c_int),
                ("learning_rate", C.c_float), ("init_std", C.c_float), ("exp_policy", C.c_int),
                ("trace_every", C.c_int), ("fast_mmd_vectors", C.c_int), ("seed", C.c_ulonglong),
                ("graph_cluster", C.c_int), ("graph_noise_regen", C.c_int)]


class GraphBatchStruct(C.Structure):
    _fields_ = [("n_graphs", C.c_int), ("n_var", C.c_int), ("n_noise_cols", C.c_int),
                ("node_col", C.POINTER(C.c_int)), ("n_parents", C.POINTER(C.c_int)), ("n_inputs", C.POINTER(C.c_int)),
                ("in_ptr", C.POINTER(C.c_int)), ("in_idx", C.POINTER(C.c_int))]


_F = C.POINTER(C.c_float)
_I = C.POINTER(C.c_int)
_V = C.c_void_p

# every symbol include/cgnn_b200.h declares: name -> (restype, argtypes)
SYMBOLS = {
    "cgnn_b200_version": (C.c_int, []),
    "cgnn_b200_create": (C.c_int, [C.POINTER(_V), C.c_int]),
    "cgnn_b200_destroy": (None, [_V]),
    "cgnn_b200_last_error": (C.c_char_p, [_V]),
    "cgnn_b200_set_stream": (C.c_int, [_V, _V]),
    "cgnn_b200_last_kernel_ms": (C.c_float, [_V]),
    "cgnn_b200_last_launch_count": (C.c_int, [_V]),
    "cgnn_b200_device_info": (C.c_int, [_V, _I, _I]),
    "cgnn_b200_last_graph_cluster": (C.c_int, [_V]),
    "cgnn_b200_last_graph_mode": (C.c_int, [_V]),
    "cgnn_b200_run_pairwise": (C.c_int, [_V, C.POINTER(Hparams), _V, _V, C.c_int, C.c_int, C.c_int,
                                         C.c_ulonglong, _V, _V, _V, _V, _V]),
    "cgnn_b200_pairwise_step_debug": (C.c_int, [_V, C.POINTER(Hparams), _V, C.c_int, C.c_int, _V, _V, _V,
                                                _V, _V, _V, _V, _V]),
    "cgnn_b200_run_graph": (C.c_int, [_V, C.POINTER(Hparams), _V, C.c_int, C.POINTER(GraphBatchStruct), C.c_int, _V, C.c_int,
                                      C.c_ulonglong, _V, _V, _V, C.c_int, _V, _V, _V]),
    "cgnn_b200_graph_step_debug": (C.c_int, [_V, C.POINTER(Hparams), _V, C.c_int, C.POINTER(GraphBatchStruct), C.c_int, _V, _V,
                                             C.c_int, _V, _V, _V, _V, _V, _V]),
    "cgnn_b200_probe_peak": (C.c_int, [_V, C.c_int, C.c_int, C.POINTER(C.c_double)]),
    "cgnn_b200_run_pairwise_noalloc": (C.c_int, [C.c_int, _V, C.POINTER(Hparams), _V, C.c_int, C.c_int, C.c_int,
                                                 C.c_ulonglong, _V, C.c_char_p, C.c_int]),
}

_lib = None


def load():
    """Load the shared library and declare every prototype.  Raises if it was not built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                "cgnn_b200: %s is missing - run `python -c 'import __graft_entry__ as g; g.build()'` "
                "(there is no CPU fallback)" % LIB_PATH)
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib
