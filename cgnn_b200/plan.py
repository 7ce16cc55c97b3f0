"""Turns a candidate graph into the plain-data *plan* the CUDA graph kernel executes.

The plan restates the model construction of the reference as data instead of TF ops:
  CGNN_tf.__init__             Code/cgnn/CGNN.py:55-90
  CGNN_confounders_tf.__init__ Code/cgnn/CGNN_confounders.py:57-107
i.e. the node build order (repeated scan of the node list, a node is emitted once all of its
parents are), each node's inputs (generated parents in get_parents order, own noise, and - hidden
variable model - one shared noise column per skeleton neighbour in get_neighbors order).

plan = dict(n_var, n_noise, nodes=[dict(col, parents=[cols], noise=[noise ids])], list_nodes, n_edges)
Column c is the c-th entry of `columns` (the order of the data matrix handed to the engine).
"""
import ctypes as C

import numpy as np

from . import _lib


def build_plan(graph, columns=None, confounders=False, canonical=False):
    """canonical=False restates the reference literally: node scan order, parent order and neighbour order are those of
    the graph object (its insertion history).  canonical=True orders everything by data column instead, so that two graph
    objects with the same edge set always give the same plan - what the search drivers need, whose cache and random
    streams are keyed by the edge set alone (the flat parameter layout, hence which initial weight a connection
    draws, follows the plan's order)."""
    if confounders:
        skel = graph.skeleton
        list_nodes = skel.get_list_nodes()                          # CGNN_confounders.py:57
        skel_edges = skel.get_list_edges_without_duplicate()        # CGNN_confounders.py:64
    else:
        list_nodes = graph.get_list_nodes()                         # CGNN.py:55
        skel_edges = []
    columns = list(columns) if columns is not None else list(list_nodes)
    col = {v: c for c, v in enumerate(columns)}
    n_var = len(list_nodes)
    if sorted(col[v] for v in list_nodes) != list(range(n_var)):
        raise ValueError("graph nodes and data columns do not match")
    if canonical:
        list_nodes = sorted(list_nodes, key=col.get)
        skel_edges = sorted((tuple(sorted((e[0], e[1]), key=col.get)) for e in skel_edges), key=lambda e: (col[e[0]], col[e[1]]))
    shared = {}
    for k, e in enumerate(skel_edges):
        shared[(e[0], e[1])] = shared[(e[1], e[0])] = k
    parents = {v: (sorted(graph.get_parents(v), key=col.get) if canonical else graph.get_parents(v)) for v in list_nodes}
    emitted, order = set(), []
    while len(order) < n_var:                                       # CGNN.py:63-69
        before = len(order)
        for v in list_nodes:
            if v not in emitted and all(p in emitted for p in parents[v]):
                emitted.add(v)
                order.append(v)
        if len(order) == before:
            raise ValueError("cyclic graph: no build order exists")
    nodes = []
    for v in order:
        noise = [col[v]]
        if confounders:
            nbs = graph.skeleton.get_neighbors(v)
            noise += [n_var + shared[(nb, v)] for nb in (sorted(nbs, key=col.get) if canonical else nbs)]
        nodes.append(dict(col=col[v], parents=[col[p] for p in parents[v]], noise=noise))
    return dict(n_var=n_var, n_noise=n_var + len(skel_edges), nodes=nodes, list_nodes=list(list_nodes),
                columns=columns, n_edges=sum(len(nd["parents"]) for nd in nodes))


def n_params(plan, h):
    return sum((len(nd["parents"]) + len(nd["noise"])) * h + 2 * h + 1 for nd in plan["nodes"])


class GraphBatch:
    """cgnn_graph_batch (include/cgnn_b200.h) for a list of plans; keeps the numpy arrays alive."""

    def __init__(self, plans):
        n_var, n_noise = plans[0]["n_var"], plans[0]["n_noise"]
        if any(p["n_var"] != n_var or p["n_noise"] != n_noise for p in plans):
            raise ValueError("all graphs of a batch must share variables and noise columns")
        self.node_col = np.array([[nd["col"] for nd in p["nodes"]] for p in plans], np.int32)
        self.n_parents = np.array([[len(nd["parents"]) for nd in p["nodes"]] for p in plans], np.int32)
        self.n_inputs = np.array([[len(nd["parents"]) + len(nd["noise"]) for nd in p["nodes"]] for p in plans], np.int32)
        idx, ptr = [], [0]
        for p in plans:
            for nd in p["nodes"]:
                idx += nd["parents"] + nd["noise"]
            ptr.append(len(idx))
        self.in_ptr = np.array(ptr, np.int32)
        self.in_idx = np.array(idx, np.int32)
        ip = lambda a: a.ctypes.data_as(C.POINTER(C.c_int))
        self.struct = _lib.GraphBatchStruct(len(plans), n_var, n_noise, ip(self.node_col), ip(self.n_parents),
                                            ip(self.n_inputs), ip(self.in_ptr), ip(self.in_idx))
