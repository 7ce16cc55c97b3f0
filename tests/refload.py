"""Loads the reference's own Python modules (read-only, /root/reference) for pinning tests.

Only available in the build container; on the GPU box /root/reference does not exist and
`available()` is False - tests then rely on the frozen fixtures in tests/golden/ alone.
The package's __init__ is bypassed (it imports TensorFlow and a broken generators module,
Code/cgnn/__init__.py:1,6); `tensorflow` is satisfied by whatever module is passed in - an empty
stub for host-logic tests, oracle.tf1_shim for arithmetic pinning.
"""
import importlib
import os
import sys
import types

REF = "/root/reference/Code/cgnn"


def available():
    return os.path.isdir(REF)


def load(tf_module=None, name="refcgnn"):
    """Returns a dict of reference modules: Graph, Formats, Settings, Loss, GNN, CGNN, CGNN_confounders, PairwiseModel."""
    import pandas as pd
    if not hasattr(pd.DataFrame, "as_matrix"):          # removed in pandas 1.0; the reference predates it
        pd.DataFrame.as_matrix = lambda self: self.values
        pd.Series.as_matrix = lambda self: self.values
    for k in [k for k in sys.modules if k == name or k.startswith(name + ".")]:
        del sys.modules[k]
    sys.modules["tensorflow"] = tf_module if tf_module is not None else types.ModuleType("tensorflow")
    pkg = types.ModuleType(name)
    pkg.__path__ = [REF]
    sys.modules[name] = pkg
    out = {}
    for m in ("utils.Graph", "utils.Formats", "utils.Settings", "utils.Loss", "PairwiseModel", "GraphModel",
              "GNN", "CGNN", "CGNN_confounders"):
        out[m.split(".")[-1]] = importlib.import_module(name + "." + m)
    return out
