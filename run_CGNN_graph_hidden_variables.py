"""CGNN with hidden common causes on a skeleton (entry point of the reference's
run_CGNN_graph_hidden_variables.py:1-33: same inputs and output files).

    python run_CGNN_graph_hidden_variables.py [data.csv skeleton.csv] [--runs 32] [--train 1000] [--test 500]

Writes <data>_printout.csv, <data>_pairwise_predictions.csv, <data>_confounders_predictions.csv.
"""
import argparse
import os

import pandas as pd

import cgnn
from cgnn_b200.utils.scaling import scale

HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("datafile", nargs="?", default=os.path.join(HERE, "examples", "Example_graph_confounders_numdata.csv"))
    ap.add_argument("skeletonfile", nargs="?", default=os.path.join(HERE, "examples", "Example_graph_confounders_skeleton.csv"))
    ap.add_argument("--runs", type=int, default=32)
    ap.add_argument("--train", type=int, default=cgnn.SETTINGS.train_epochs)
    ap.add_argument("--test", type=int, default=cgnn.SETTINGS.test_epochs)
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()

    cgnn.SETTINGS.GPU = True
    cgnn.SETTINGS.NB_GPU = a.gpus
    cgnn.SETTINGS.NB_JOBS = 8
    cgnn.SETTINGS.NB_RUNS = a.runs
    cgnn.SETTINGS.train_epochs, cgnn.SETTINGS.test_epochs = a.train, a.test

    prefix = a.out or a.datafile
    data = pd.read_csv(a.datafile)
    skeleton = cgnn.UndirectedGraph(pd.read_csv(a.skeletonfile))
    data = pd.DataFrame(scale(data.values), columns=data.columns)      # run_CGNN_graph_hidden_variables.py:20

    gnn = cgnn.GNN(backend="TensorFlow")
    p_directed_graph = gnn.orient_graph_confounders(data, skeleton, printout=prefix + "_printout.csv")
    gnn_res = pd.DataFrame(p_directed_graph.get_list_edges(descending=True), columns=["Cause", "Effect", "Score"])
    gnn_res.to_csv(prefix + "_pairwise_predictions.csv")

    model = cgnn.CGNN_confounders(backend="TensorFlow")
    directed_graph = model.orient_directed_graph(data, p_directed_graph)
    cgnn_res = pd.DataFrame(directed_graph.get_list_edges(descending=True), columns=["Cause", "Effect", "Score"])
    cgnn_res.to_csv(prefix + "_confounders_predictions.csv")
    return gnn_res, cgnn_res


if __name__ == "__main__":
    main()
