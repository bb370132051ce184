# -*- coding: utf-8 -*-
"""
Per-track feature files.  The reference stores one deepdish-HDF5 dict per track
(`dd.io.load(path)` / `dd.io.save`, acoss/algorithms/algorithm_template.py:90,193;
schema README.md:94-114).  deepdish / h5py / pytables are not installed in this
image, so `load`/`save` use deepdish when importable and otherwise a pickled dict
with the same keys under the same `.h5` path (the synthetic generator writes that).
"""
import pickle

import numpy as np

try:  # pragma: no cover - not available in this image
    import deepdish as _dd
    if not hasattr(_dd, "io") or getattr(_dd.io, "__name__", "") == "_DeepdishIO":
        _dd = None
except Exception:  # ImportError and friends
    _dd = None

_HDF5_MAGIC = b"\x89HDF\r\n\x1a\n"


def load(path):
    with open(path, "rb") as f:
        head = f.read(8)
    if head == _HDF5_MAGIC:
        if _dd is None:
            raise IOError("%s is an HDF5 file but deepdish is not installed" % path)
        return _dd.io.load(path)
    with open(path, "rb") as f:
        return pickle.load(f)


def save(path, obj):
    if isinstance(obj, dict):
        obj = {k: (np.array(v) if isinstance(v, np.memmap) else v) for k, v in obj.items()}
    with open(path, "wb") as f:
        pickle.dump(obj, f, protocol=4)
