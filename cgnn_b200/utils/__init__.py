from .Settings import SETTINGS
from .Formats import CCEPC_PairsFileReader
from .Graph import DirectedGraph, UndirectedGraph
