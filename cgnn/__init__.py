"""`import cgnn` compatibility shim: the reference's package name, served by cgnn_b200."""
from cgnn_b200 import *                      # noqa: F401,F403
from cgnn_b200 import SETTINGS, GNN, CGNN, CGNN_confounders, DirectedGraph, UndirectedGraph, utils  # noqa: F401
