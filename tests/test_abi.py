"""The C-ABI shared library builds, loads and exports every symbol include/cgnn_b200.h declares
(no compute calls here: there is no GPU in the CPU test run)."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "cgnn_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(cgnn_b200_\w+)\s*\(", text)))


def test_library_builds_and_exports_every_declared_symbol():
    import __graft_entry__ as ge
    ge.build()
    from cgnn_b200 import _lib
    lib = ctypes.CDLL(_lib.LIB_PATH)
    names = declared_symbols()
    assert len(names) >= 14
    for n in names:
        assert hasattr(lib, n), "libcgnn_b200.so does not export %s" % n
    assert sorted(_lib.SYMBOLS) == names, "cgnn_b200/_lib.py must bind exactly the declared entry points"
    assert lib.cgnn_b200_version() == 100


def test_hparams_struct_layout_matches_header():
    from cgnn_b200 import _lib
    assert ctypes.sizeof(_lib.Hparams) == 48 and _lib.Hparams.seed.offset == 32 and _lib.Hparams.graph_cluster.offset == 40
    assert ctypes.sizeof(_lib.GraphBatchStruct) == 16 + 5 * 8


def test_header_constants_match_the_host_layer():
    """Limits and the automatic exp policy: the numbers in the header are the numbers the Python layer uses."""
    text = open(os.path.join(ROOT, "include", "cgnn_b200.h")).read()
    defs = {k: int(v) for k, v in re.findall(r"#define\s+(CGNN_\w+)\s+\(?(-?\d+)\)?\s", text)}
    from cgnn_b200 import SETTINGS
    assert defs["CGNN_EXP_POLICY_AUTO"] == -1 and SETTINGS.exp_policy == defs["CGNN_EXP_POLICY_AUTO"]
    assert defs["CGNN_EXP_POLICY_PAIRWISE"] == 3 and defs["CGNN_EXP_POLICY_GRAPH"] == 1
    assert defs["CGNN_MAX_ROWS_PAIRWISE"] >= SETTINGS.max_nb_points and defs["CGNN_MAX_ROWS_GRAPH"] >= SETTINGS.max_nb_points
    assert defs["CGNN_MAX_FAST_VECTORS"] >= SETTINGS.nb_vectors_approx_MMD


def test_no_gpu_means_loud_failure():
    """Without an sm_100 device the product raises; it never falls back to the oracle or the CPU."""
    import torch
    if torch.cuda.is_available():
        return
    import pytest
    from cgnn_b200 import Engine, EngineError
    with pytest.raises(EngineError):
        Engine(0)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "cgnn_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), "%s imports oracle" % f
