"""Base class of graph-level causal models; same dispatch as the reference's GraphModel
(Code/cgnn/GraphModel.py:9-47): `predict(df_data, graph=None)` routes on the type of `graph`."""
from .utils.Graph import DirectedGraph, UndirectedGraph


class GraphModel(object):
    def __init__(self):
        super(GraphModel, self).__init__()

    def predict(self, df_data, graph=None, **kwargs):
        if graph is None:
            return self.create_graph_from_data(df_data, **kwargs)
        if type(graph) == DirectedGraph:
            return self.orient_directed_graph(df_data, graph, **kwargs)
        if type(graph) == UndirectedGraph:
            return self.orient_undirected_graph(df_data, graph, **kwargs)
        print("Unknown Graph type")
        raise ValueError

    def orient_undirected_graph(self, data, umg, **kwargs):
        raise NotImplementedError

    def orient_directed_graph(self, data, dag, **kwargs):
        raise NotImplementedError

    def create_graph_from_data(self, data, **kwargs):
        raise NotImplementedError
