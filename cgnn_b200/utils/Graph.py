"""Graph containers with the reference's API and ordering semantics.

API-compatible with /root/reference/Code/cgnn/utils/Graph.py (DirectedGraph, UndirectedGraph):
the hill-climbing drivers depend on *orderings* that fall out of that file's dict-of-dict
storage - node order (Graph.py:92-107), parent order (:79-90), edge order by (weight, [cause,
effect]) (:109-138), child-list order inside get_dict_nw (:160-174, used for the tabu test) and the
fact that re-adding a reversed edge appends its new cause at the END of the node order when it
had no outgoing edge before (:269-292).  The storage here is therefore the same shape - an
insertion-ordered ``{cause: {effect: weight}}`` - with independent code; tests/test_graph_utils.py
pins every method against the reference file (which runs unmodified in this container) and
against frozen fixtures.

Deliberately preserved behaviour of live paths: remove_cycles always deletes the
weakest edge of the first cycle (the reference tests ``self.cycles()`` instead of the trial
graph's, Graph.py:311-315, so its "reverse" branch is unreachable); reverse_edge treats a weight
of 0 as "keep the old weight" (:279-280).
"""
import numpy as np


class Graph(object):
    """Ordered adjacency ``{cause: {effect: weight}}``; subclasses define edge insertion/removal."""

    def __init__(self, df=None, adjacency_matrix=False):
        self._graph = {}
        if df is None:
            return
        if adjacency_matrix:                                   # Graph.py:37-43
            m = np.asarray(df.values)
            names = list(df.columns)
            links = [(names[i], names[j], m[i, j]) for i in range(m.shape[0]) for j in range(m.shape[0])
                     if i != j and m[i, j] > 0.001]
        else:                                                  # one edge per row: (node1, node2[, weight])
            links = [tuple(row) for row in df.itertuples(index=False, name=None)]
        self.add_multiple_edges(links)

    # -- construction ---------------------------------------------------------------------------
    def add_multiple_edges(self, connections):
        for link in connections:
            self.add(*link[:3])

    def add(self, node1, node2, weight=1):
        raise NotImplementedError

    def remove_edge(self, node1, node2):
        raise NotImplementedError

    def _out(self, node):
        return self._graph.get(node, {})

    # -- queries --------------------------------------------------------------------------------
    def get_parents(self, node):
        """Causes of `node`, in node-insertion order of the causes."""
        return [src for src, targets in self._graph.items() if node in targets]

    def get_list_nodes(self):
        """Every cause in insertion order, each followed by its not-yet-seen effects."""
        seen = {}
        for src, targets in self._graph.items():
            seen.setdefault(src)
            for dst in targets:
                seen.setdefault(dst)
        return list(seen)

    def get_list_edges(self, order_by_weight=True, descending=False, return_weights=True):
        """[[cause, effect, weight], ...]; ordered by (weight, [cause, effect]) unless told otherwise."""
        rows = [(w, [src, dst]) for src, targets in self._graph.items() for dst, w in targets.items()]
        if order_by_weight:
            rows.sort(reverse=bool(descending))
        if return_weights:
            return [[e[0], e[1], w] for w, e in rows]
        return [e for _, e in rows]

    def get_adjacency_matrix(self):
        nodes = self.get_list_nodes()
        pos = {n: k for k, n in enumerate(nodes)}
        m = np.zeros((len(nodes), len(nodes)))
        for src, targets in self._graph.items():
            for dst, w in targets.items():
                m[pos[src], pos[dst]] = w
        return m, nodes

    def get_dict_nw(self):
        """Weight-free structure {node: [effects...]}, childless effects included with []."""
        out = {}
        for src, targets in self._graph.items():
            for dst in targets:
                out.setdefault(src, []).append(dst)
                out.setdefault(dst, [])
        return out

    def remove_node(self, node):
        for src in list(self._graph):
            self._graph[src].pop(node, None)
        self._graph.pop(node, None)

    def __str__(self):
        return "{}({})".format(type(self).__name__, self._graph)


class DirectedGraph(Graph):
    def __init__(self, df=None, adjacency_matrix=False, skeleton=False):
        self.skeleton = skeleton
        super(DirectedGraph, self).__init__(df, adjacency_matrix)

    def add(self, node1, node2, weight=1):
        self._graph.setdefault(node1, {})[node2] = weight
        return self

    def remove_edge(self, node1, node2):
        del self._graph[node1][node2]
        if not self._graph[node1]:
            del self._graph[node1]

    def set_weight(self, node1, node2, weight):
        self._graph.setdefault(node1, {})[node2] = weight

    def reverse_edge(self, node1, node2, weight=None):
        w = weight if weight else self._graph[node1][node2]
        self.remove_edge(node1, node2)
        self.add(node2, node1, w)

    def is_cyclic(self):
        """Depth-first search for a back edge over the weight-free structure."""
        succ = self.get_dict_nw()
        state = {}                      # node -> 1 on the current path, 2 finished

        def descend(v):
            state[v] = 1
            for nxt in succ.get(v, ()):
                if state.get(nxt) == 1 or (nxt not in state and descend(nxt)):
                    return True
            state[v] = 2
            return False

        return any(v not in state and descend(v) for v in succ)

    def cycles(self):
        """All simple cycles as node lists [start, ..., start], enumerated from every start node with a
        LIFO frontier (the enumeration order matters to remove_cycles, which uses the first one)."""
        succ = self.get_dict_nw()
        found = []
        for start in succ:
            frontier = [(start, [])]
            while frontier:
                node, path = frontier.pop()
                if path and node == start:
                    found.append([start] + path)
                    continue
                for nxt in succ[node]:
                    if nxt not in path:
                        frontier.append((nxt, path + [nxt]))
        return found

    def remove_cycles(self, verbose=True):
        """Delete, cycle after cycle, the lowest-weight edge of the first enumerated cycle
        (behaviour of Graph.py:294-322 as it actually runs; see module docstring)."""
        by_weight = self.get_list_edges(return_weights=False)
        while self.is_cyclic():
            first = self.cycles()[0]
            hops = [first[k:k + 2] for k in range(len(first) - 1)]
            victim = next(e for e in by_weight if e in hops)
            self.remove_edge(victim[0], victim[1])
            if verbose:
                print("Link {} got deleted !".format(victim))

    def remove_cycle_without_deletion(self):
        """Walk a snapshot of the structure depth-first and reverse every edge that closes a path
        (Graph.py:324-351).  The snapshot is not refreshed while edges flip, as in the reference."""
        succ = self.get_dict_nw()
        on_path, done = set(), set()

        def walk(v):
            if v in done:
                return
            done.add(v)
            on_path.add(v)
            for nxt in succ.get(v, ()):
                if nxt in on_path:
                    self.reverse_edge(v, nxt)
                else:
                    walk(nxt)
            on_path.discard(v)

        for v in succ:
            walk(v)

    def get_correlation_matrix(self, sigma):
        nodes = self.skeleton.get_list_nodes()
        pos = {n: k for k, n in enumerate(nodes)}
        m = np.eye(len(nodes))
        for a, b in self.skeleton.get_list_edges_without_duplicate():
            m[pos[a], pos[b]] = m[pos[b], pos[a]] = sigma
        return m


class UndirectedGraph(Graph):
    def add(self, node1, node2, weight=1):
        self._graph.setdefault(node1, {})[node2] = weight
        self._graph.setdefault(node2, {})[node1] = weight

    def remove_edge(self, node1, node2):
        del self._graph[node1][node2]
        del self._graph[node2][node1]
        for n in (node1, node2):
            if n in self._graph and not self._graph[n]:
                del self._graph[n]

    def get_neighbors(self, node):
        return self.get_parents(node)

    def get_list_edges_without_duplicate(self):
        """Each undirected edge once, as first met when scanning causes then effects."""
        out = []
        for src, targets in self._graph.items():
            for dst in targets:
                if [dst, src] not in out:
                    out.append([src, dst])
        return out
