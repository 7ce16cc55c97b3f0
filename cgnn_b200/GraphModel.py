"""Common front door of the graph-level models (CGNN, CGNN_confounders).

`predict(df_data, graph=None)` keeps the behaviour of the reference's base class
(Code/cgnn/GraphModel.py:9-47): which method runs depends on the exact type of `graph` -
nothing -> build a graph from raw data, a DirectedGraph -> improve its orientation, an
UndirectedGraph -> orient it first; any other object is reported and refused with ValueError.
"""
from .utils.Graph import DirectedGraph, UndirectedGraph


class GraphModel(object):
    # exact type -> name of the method that handles it (subclasses of the graph classes are not accepted,
    # as in the reference, which compares with type(...) ==)
    _ROUTES = {type(None): "create_graph_from_data",
               DirectedGraph: "orient_directed_graph",
               UndirectedGraph: "orient_undirected_graph"}

    def __init__(self):
        super(GraphModel, self).__init__()

    def predict(self, df_data, graph=None, **kwargs):
        handler = self._ROUTES.get(type(graph))
        if handler is None:
            print("Unknown Graph type")
            raise ValueError
        method = getattr(self, handler)
        return method(df_data, **kwargs) if graph is None else method(df_data, graph, **kwargs)

    # the three operations a concrete model provides
    def create_graph_from_data(self, data, **kwargs):
        raise NotImplementedError

    def orient_directed_graph(self, data, dag, **kwargs):
        raise NotImplementedError

    def orient_undirected_graph(self, data, umg, **kwargs):
        raise NotImplementedError
