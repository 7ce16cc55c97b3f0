"""cgnn_b200 - B200-native engine for the CGNN hot path.

Public surface = the reference's (Code/cgnn/__init__.py:1-11): SETTINGS, GNN, CGNN,
CGNN_confounders, DirectedGraph, UndirectedGraph, utils.CCEPC_PairsFileReader - plus Engine, the
handle on the CUDA library.  `import cgnn` (the shim package at the repository root) gives the same
names, so the reference's run scripts work unchanged.
"""
from .utils.Settings import SETTINGS
from .utils.Graph import DirectedGraph, UndirectedGraph
from .engine import Engine, EngineError
from .GNN import GNN
from .CGNN import CGNN
from .CGNN_confounders import CGNN_confounders
from . import utils

__all__ = ["DirectedGraph", "UndirectedGraph", "CGNN", "CGNN_confounders", "GNN", "SETTINGS", "Engine", "EngineError"]
