"""Text model dump / load, byte-compatible with the reference.

ref: BCDServer::saveModel (src/learner/bcd.h:251-272), KVStore::writeToFile
(src/parameter/kv_store.h:63-73), ModelEvaluation::run (model_evaluation.h:22-33).
One `"<uint64 key>\\t<value>\\n"` per non-zero, non-NaN weight, default ostream formatting
(6 significant digits == "%g").
"""
from __future__ import annotations

import glob
import os
import re

import numpy as np


def write_model_text(path: str, keys: np.ndarray, values: np.ndarray) -> str:
    d = os.path.dirname(path)
    if d and not os.path.isdir(d):
        os.makedirs(d, exist_ok=True)
    keep = (values != 0) & ~np.isnan(values)
    with open(path, "w") as f:
        for k, v in zip(keys[keep].tolist(), values[keep].tolist()):
            f.write("%d\t%g\n" % (k, v))
    return path


def search_files(pattern: str) -> list:
    """Files whose basename matches the regex in `pattern` (ref data/common.cc:100-150)."""
    d, base = os.path.split(pattern)
    d = d or "."
    if not os.path.isdir(d):
        return []
    try:
        rx = re.compile(base)
    except re.error:
        rx = None
    out = []
    for name in sorted(os.listdir(d)):
        if (rx is not None and rx.fullmatch(name)) or name == base or \
                (rx is not None and rx.match(name) and name.startswith(base.rstrip(".*"))):
            out.append(os.path.join(d, name))
    if not out:
        out = sorted(glob.glob(pattern + "*"))
    return out


def read_model_text(pattern_or_files) -> dict:
    """Load `key value` pairs from every file matching the pattern into a dict."""
    files = pattern_or_files if isinstance(pattern_or_files, (list, tuple)) else \
        search_files(pattern_or_files)
    model = {}
    for fn in files:
        with open(fn) as f:
            toks = f.read().split()
        for i in range(0, len(toks) - 1, 2):
            model[int(toks[i])] = float(toks[i + 1])
    return model
