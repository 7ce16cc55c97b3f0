"""Pairwise GNN on a cause-effect-pairs CSV (entry point of the reference's
run_GNN_pairwise_inference.py:1-28: same settings, same input format, same output files).

    python run_GNN_pairwise_inference.py [pairs.csv] [--runs 32] [--train 1000] [--test 500]

Writes <pairs.csv>_printout.csv (SampleID, Predictions) and <pairs.csv>_predictions_GNN.csv.
"""
import argparse
import os

import pandas as pd

import cgnn
from cgnn.utils import CCEPC_PairsFileReader as CC

HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("datafile", nargs="?", default=os.path.join(HERE, "examples", "Example_pairwise_pairs.csv"))
    ap.add_argument("--runs", type=int, default=32)          # run_GNN_pairwise_inference.py:17
    ap.add_argument("--train", type=int, default=cgnn.SETTINGS.train_epochs)
    ap.add_argument("--test", type=int, default=cgnn.SETTINGS.test_epochs)
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--out", default=None, help="output prefix (default: next to the data file)")
    a = ap.parse_args()

    cgnn.SETTINGS.GPU = True
    cgnn.SETTINGS.NB_GPU = a.gpus
    cgnn.SETTINGS.NB_JOBS = 8                                 # accepted, unused: one launch scores the batch
    cgnn.SETTINGS.h_layer_dim = 30                            # run_GNN_pairwise_inference.py:9
    cgnn.SETTINGS.use_Fast_MMD = False
    cgnn.SETTINGS.NB_RUNS = a.runs
    cgnn.SETTINGS.train_epochs, cgnn.SETTINGS.test_epochs = a.train, a.test

    prefix = a.out or a.datafile
    print("Processing " + a.datafile + "...")
    data = CC(a.datafile, scale=True)
    model = cgnn.GNN(backend="TensorFlow")                    # the reference's backend name is an alias here
    predictions = model.predict_dataset(data, printout=prefix + "_printout.csv")
    predictions = pd.DataFrame(predictions, columns=["Predictions"])
    print("Processed " + a.datafile)
    predictions.to_csv(prefix + "_predictions_GNN.csv")
    return predictions


if __name__ == "__main__":
    main()
