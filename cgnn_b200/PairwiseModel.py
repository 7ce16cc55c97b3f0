"""Base class of pairwise causal-direction models, API-compatible with the reference's
`Pairwise_Model` (Code/cgnn/PairwiseModel.py:11-125): `predict_proba`, `predict_dataset`,
`orient_graph`, `orient_graph_confounders` with the same arguments, return values and printout
files (columns ``SampleID, Predictions``).

Difference in execution only: the reference walks pairs / skeleton edges one by one and scores each
with `predict_proba`; here every pair (or edge) of the call is collected first and handed to
`predict_proba_batch`, which a subclass implements as ONE engine submission.  The default
`predict_proba_batch` falls back to a loop over `predict_proba`, so third-party subclasses that only
define `predict_proba` keep working.
"""
import numpy as np
from pandas import DataFrame

from .utils.Graph import DirectedGraph
from .utils.scaling import scale


def _column(values):
    return scale(np.asarray(values, dtype=np.float64).reshape(-1, 1))


class Pairwise_Model(object):
    def __init__(self):
        super(Pairwise_Model, self).__init__()

    def predict_proba(self, a, b, idx=0, **kwargs):
        """Score of one pair: > 0 if a causes b, < 0 if b causes a (PairwiseModel.py:21-30)."""
        raise NotImplementedError

    def predict_proba_batch(self, pairs, idx0=0, **kwargs):
        """pairs: list of (a, b) column arrays.  Returns the list of scores."""
        return [self.predict_proba(a, b, idx0 + k, **kwargs) for k, (a, b) in enumerate(pairs)]

    @staticmethod
    def _printout(path, names, scores):
        if path is not None:
            DataFrame(list(zip(names, scores)), columns=["SampleID", "Predictions"]).to_csv(path, index=False)

    def predict_dataset(self, x, printout=None, **kwargs):
        """Scores every row of a CEPC DataFrame (SampleID, A, B); each column is standardised first
        (PairwiseModel.py:32-55).  Returns the list of predictions in row order."""
        pairs = [(_column(a), _column(b)) for a, b in zip(x["A"], x["B"])]
        pred = list(self.predict_proba_batch(pairs, 0, **kwargs))
        self._printout(printout, list(x["SampleID"]), pred)
        return pred

    def _orient(self, df_data, umg, graph, printout, **kwargs):
        edges = umg.get_list_edges_without_duplicate()
        pairs = [(_column(df_data[a].values), _column(df_data[b].values)) for a, b in edges]
        weights = list(self.predict_proba_batch(pairs, 0, **kwargs))
        for (a, b), w in zip(edges, weights):
            if w > 0:                       # a causes b (PairwiseModel.py:77-80)
                graph.add(a, b, w)
            else:
                graph.add(b, a, abs(w))
        self._printout(printout, ["%s-%s" % (a, b) for a, b in edges], weights)
        return graph

    def orient_graph(self, df_data, umg, printout=None, **kwargs):
        """Orients every edge of an UndirectedGraph with the pairwise score, then reverses the edges
        that close a cycle (PairwiseModel.py:57-90).  Returns a weighted DirectedGraph."""
        graph = self._orient(df_data, umg, DirectedGraph(), printout, **kwargs)
        graph.remove_cycle_without_deletion()
        return graph

    def orient_graph_confounders(self, df_data, umg, printout=None, **kwargs):
        """Same for the hidden-variable model: the result keeps `umg` as its skeleton and cycles are
        removed by deleting edges (PairwiseModel.py:92-125)."""
        graph = self._orient(df_data, umg, DirectedGraph(skeleton=umg), printout, **kwargs)
        graph.remove_cycles()
        return graph
