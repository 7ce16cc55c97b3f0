"""Column standardisation with the semantics of sklearn.preprocessing.scale, which the reference
applies before every pairwise call (PairwiseModel.py:46-47,75,111; Formats.py:44-46; CGNN.py:365):
zero mean, unit POPULATION standard deviation (ddof=0); a constant column is only centred."""
import numpy as np


def scale(x):
    a = np.asarray(x, dtype=np.float64)
    mean = a.mean(axis=0)
    std = a.std(axis=0)
    std = np.where(std == 0.0, 1.0, std)
    out = (a - mean) / std
    return out - out.mean(axis=0)          # sklearn re-centres once to remove the rounding residue
