"""Synthetic layered-DAG data with the recipe of the reference's generator package, in working order.

The reference's `RandomGraphGenerator` (Code/cgnn/generators/random_graph_generator.py:25-189) is not
importable today (`sklearn.mixture.GMM`, functions_default.py:14) and its `generate_pairs` calls a
method that does not exist (:162).  This is a re-implementation of what it describes, on a seeded
numpy Generator: uniform root causes (functions_default.py:24), each child = standardised combination
of cubic-spline mechanisms of its parents (:33-41) plus additive or multiplicative noise (:28-29,
random_graph_generator.py:87-101), nodes added layer by layer with up to `max_joint_causes` parents
drawn from earlier layers (:63-76); optional binning of a share of the variables (:113-129).
Same constructor arguments, `generate()`, `get_data()`, `save_data()`, `generate_pairs()`.
It is host-side data preparation for BASELINE configs[3]/[4] - not part of the GPU hot path.
"""
import numpy as np
import pandas as pd

from ..synth import _spline_mechanism, _standardise
from ..utils.Graph import DirectedGraph


def series_to_cepc(a, b, idx):
    """One row of the ChaLearn cause-effect-pairs format (space-separated floats with a leading blank)."""
    fmt = lambda v: " " + " ".join(repr(float(x)) for x in np.asarray(v))
    return ["pair%d" % idx, fmt(a), fmt(b)]


class RandomGraphGenerator(object):
    def __init__(self, num_nodes=200, max_joint_causes=4, noise_qty=.7, number_points=500, categorical_rate=.20, seed=0):
        self.nodes, self.num_max_parents = int(num_nodes), int(max_joint_causes)
        self.noise, self.n_points, self.cat_rate = float(noise_qty), int(number_points), float(categorical_rate)
        self.rng = np.random.default_rng(seed)
        self.graph = self.data = self.result_links = self.cat_data = self.cat_var = None

    def _noise(self):
        return self.noise * self.rng.uniform() * self.rng.normal(size=self.n_points)

    def generate(self, gen_cat=True):
        rng = self.rng
        n_roots = int(rng.integers(2, max(3, int(self.nodes / np.floor(np.sqrt(self.nodes))))))
        n_roots = min(n_roots, self.nodes)
        cols = {"V%d" % i: rng.uniform(-1.0, 1.0, self.n_points) for i in range(n_roots)}
        layers, links, made = [list(range(n_roots))], [], n_roots
        while made < self.nodes:
            width = int(rng.integers(2, len(layers[-1]) + 2))
            layer = []
            earlier = [v for l in layers for v in l]
            for _ in range(min(width, self.nodes - made)):
                parents = sorted(set(int(rng.choice(earlier)) for _ in range(self.num_max_parents)))
                parts = [_standardise(_spline_mechanism(cols["V%d" % p], rng)) for p in parents]
                links += [["V%d" % p, "V%d" % made] for p in parents]
                total = np.sum(parts, axis=0)
                if rng.integers(0, 2):                                   # multiplicative noise
                    nz = self._noise()
                    total = (total + abs(total.min())) * (nz + abs(nz.min()))
                else:
                    total = total + self._noise()
                cols["V%d" % made] = _standardise(total)
                layer.append(made)
                made += 1
            layers.append(layer)
        self.data = pd.DataFrame(cols)
        self.result_links = pd.DataFrame(links, columns=["Cause", "Effect"])
        if gen_cat:
            k = int(np.ceil(self.cat_rate * self.nodes))
            chosen = sorted(int(v) for v in rng.choice(self.nodes, size=k, replace=False))
            self.cat_data = self.data.copy()
            for v in chosen:
                n_bins = int(rng.integers(2, 20))
                edges = np.linspace(-3, 3, n_bins + 1)
                self.cat_data["V%d" % v] = np.digitize(_standardise(self.data["V%d" % v].values), edges) - 1
            self.cat_var = pd.DataFrame(chosen)
        self.graph = DirectedGraph()
        self.graph.add_multiple_edges([l + [1] for l in links])
        return self.get_data()

    def get_data(self):
        if self.graph is None:
            print("Please compute graph using .generate(), graph not build yet")
            raise NameError
        return self.graph, self.data, self.cat_data, self.cat_var

    def save_data(self, filename):
        self.get_data()
        self.result_links.to_csv(filename + "_target.csv", sep=",", index=False)
        self.data.to_csv(filename + "_numdata.csv", sep=",", index=False)
        if self.cat_data is not None:
            self.cat_data.to_csv(filename + "_catdata.csv", sep=",", index=False)
            self.cat_var.to_csv(filename + "_catindex.csv", sep=",", index=False)

    def generate_pairs(self, num_pairs):
        """CEPC-format pairs drawn from the edges of freshly generated graphs; target +1 if A causes B."""
        pairs, targets = [], []
        while len(pairs) < num_pairs:
            self.generate(gen_cat=False)
            for cause, effect in self.result_links.values.tolist():
                if len(pairs) >= num_pairs:
                    break
                k = len(pairs)
                if self.rng.integers(0, 2):
                    pairs.append(series_to_cepc(self.data[cause], self.data[effect], k)); targets.append(["pair%d" % k, 1.0])
                else:
                    pairs.append(series_to_cepc(self.data[effect], self.data[cause], k)); targets.append(["pair%d" % k, -1.0])
        return (pd.DataFrame(pairs, columns=["SampleID", "A", "B"]), pd.DataFrame(targets, columns=["SampleID", "Target"]))
