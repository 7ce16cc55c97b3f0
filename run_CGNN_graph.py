"""Skeleton -> DAG with pairwise GNN, then CGNN hill climbing (entry point of the reference's
run_CGNN_graph.py:1-31: same inputs and output files).

    python run_CGNN_graph.py [data.csv skeleton.csv] [--runs 32] [--train 1000] [--test 500]

Writes <data>_printout.csv, <data>_pairwise_predictions.csv and <data>_predictions.csv
(columns Cause, Effect, Score; descending score).
"""
import argparse
import os

import pandas as pd

import cgnn

HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("datafile", nargs="?", default=os.path.join(HERE, "examples", "Example_graph_numdata.csv"))
    ap.add_argument("skeletonfile", nargs="?", default=os.path.join(HERE, "examples", "Example_graph_skeleton.csv"))
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
    print("Processing " + a.datafile + "...")
    umg = cgnn.UndirectedGraph(pd.read_csv(a.skeletonfile))
    data = pd.read_csv(a.datafile)

    gnn = cgnn.GNN(backend="TensorFlow")
    p_directed_graph = gnn.orient_graph(data, umg, printout=prefix + "_printout.csv")
    gnn_res = pd.DataFrame(p_directed_graph.get_list_edges(descending=True), columns=["Cause", "Effect", "Score"])
    gnn_res.to_csv(prefix + "_pairwise_predictions.csv")

    model = cgnn.CGNN(backend="TensorFlow")
    directed_graph = model.orient_directed_graph(data, p_directed_graph)
    cgnn_res = pd.DataFrame(directed_graph.get_list_edges(descending=True), columns=["Cause", "Effect", "Score"])
    cgnn_res.to_csv(prefix + "_predictions.csv")
    print("Processed " + a.datafile)
    return gnn_res, cgnn_res


if __name__ == "__main__":
    main()
