from cgnn_b200.utils import CCEPC_PairsFileReader, SETTINGS, DirectedGraph, UndirectedGraph  # noqa: F401
