"""Readers for the reference's on-disk formats.

CCEPC_PairsFileReader keeps the name, arguments and return value of
Code/cgnn/utils/Formats.py:12-53: a ChaLearn cause-effect-pairs CSV (``SampleID,A,B`` with A and B
space-separated floats, usually with a leading space) becomes a DataFrame with columns
``SampleID, A, B`` where A and B are numpy arrays, standardised when ``scale=True``.
The cells are parsed with one vectorised numpy call per cell instead of a Python float() per number.
"""
import numpy as np
from pandas import DataFrame, read_csv

from .scaling import scale as _standardise


def _parse_cell(cell):
    return np.array(str(cell).split(), dtype=np.float64)      # split() drops leading/trailing/duplicate blanks


def CCEPC_PairsFileReader(filename, scale=True):
    raw = read_csv(filename)
    rows = []
    for sample_id, a_txt, b_txt in zip(raw["SampleID"], raw["A"], raw["B"]):
        a, b = _parse_cell(a_txt), _parse_cell(b_txt)
        if scale:
            a, b = _standardise(a), _standardise(b)
        rows.append((sample_id, a, b))
    return DataFrame(rows, columns=["SampleID", "A", "B"])


def pairs_to_batch(df, max_points=None, rng=None):
    """Stack the pairs of a CEPC DataFrame into the engine's batch layout:
    (xy float32 [n_pairs, ld_rows, 2], n_rows int32 [n_pairs])."""
    n = [len(a) for a in df["A"]]
    ld = max(n)
    xy = np.zeros((len(n), ld, 2), np.float32)
    for k, (a, b) in enumerate(zip(df["A"], df["B"])):
        xy[k, :n[k], 0] = a
        xy[k, :n[k], 1] = b
    return xy, np.asarray(n, np.int32)
