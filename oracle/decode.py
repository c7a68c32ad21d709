"""Oracle: keras decode_predictions(top=5) and the reference's JSON dump.

Follows /root/reference/models.py:42 / :67 (decode_predictions) and
/root/reference/models.py:109-126 (NpEncoder, dump_to_file).  Ties: the
reference's ``pred.argsort()[-5:][::-1]`` uses numpy's default (unstable) sort;
oracle and kernel both define "larger class index first" (a stable ascending
argsort, reversed).  Test infrastructure only.
"""
from __future__ import annotations

import json

import numpy as np


def top5_indices(probs_row: np.ndarray) -> np.ndarray:
    order = np.argsort(probs_row, kind="stable")
    return order[-5:][::-1]


def decode_predictions(probs: np.ndarray, class_index, top: int = 5):
    """probs [N,1000] -> [[(wnid, name, np.float32 prob) x top] x N]."""
    if probs.ndim != 2 or probs.shape[1] != 1000:
        raise ValueError("`decode_predictions` expects a batch of predictions (N, 1000)")
    out = []
    for row in probs:
        idx = np.argsort(row, kind="stable")[-top:][::-1]
        res = [tuple(class_index[int(i)]) + (row[i],) for i in idx]
        res.sort(key=lambda r: r[2], reverse=True)  # stable, as in Keras
        out.append(res)
    return out


class NpEncoder(json.JSONEncoder):
    def default(self, obj):
        if isinstance(obj, np.integer):
            return int(obj)
        if isinstance(obj, np.floating):
            return float(obj)
        if isinstance(obj, np.ndarray):
            return obj.tolist()
        return json.JSONEncoder.default(self, obj)


def dumps(d) -> str:
    """Text that dump_to_file writes (print adds the trailing newline)."""
    return json.dumps(d, indent=4, cls=NpEncoder) + "\n"
