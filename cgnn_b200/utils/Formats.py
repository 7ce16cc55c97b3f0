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
