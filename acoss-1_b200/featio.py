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


# ---- packed store ---------------------------------------------------
# One file per (dataset, feature): every track's (T_t, dim) float32 frames back to back, memory-mappable and in
# the exact layout acoss_upload_tracks / acoss_median_downsample take (frames + row offsets), so that loading a
# 1000- or 15 000-track set is one mmap instead of one deepdish/pickle load per track (the reference re-opens a
# per-track HDF5 file in every load_features call, algorithm_template.py:90).
#
#   bytes 0..63   header: magic "ACOSSPK1", int64 n_tracks, int64 dim, int64 total_rows, int64 labels_bytes, 24 B pad
#   then          int64 offsets[n_tracks + 1]
#   then          labels: UTF-8, '\n'-joined (the clique of every track, in dataset order), padded to 64 B
#   then          float32 frames[total_rows][dim]   (64-byte aligned)
_PK_MAGIC = b"ACOSSPK1"


def _pad64(n):
    return (n + 63) // 64 * 64


def save_packed(path, tracks, labels):
    """tracks: list of (T_t, dim) arrays; labels: list of str (one per track)."""
    tracks = [np.ascontiguousarray(t, dtype=np.float32) for t in tracks]
    if len(tracks) != len(labels) or not tracks:
        raise ValueError("save_packed: need one label per track and at least one track")
    dim = int(tracks[0].shape[1])
    if any(t.ndim != 2 or t.shape[1] != dim for t in tracks):
        raise ValueError("save_packed: every track must be (T, %d)" % dim)
    if any("\n" in str(l) for l in labels):
        raise ValueError("save_packed: labels must not contain newlines")
    offs = np.zeros(len(tracks) + 1, np.int64)
    offs[1:] = np.cumsum([t.shape[0] for t in tracks])
    lab = "\n".join(str(l) for l in labels).encode("utf-8")
    head = _PK_MAGIC + np.array([len(tracks), dim, offs[-1], len(lab)], np.int64).tobytes()
    with open(path, "wb") as f:
        f.write(head.ljust(64, b"\0"))
        f.write(offs.tobytes())
        pos = 64 + offs.nbytes
        f.write(b"\0" * (_pad64(pos) - pos))
        f.write(lab.ljust(_pad64(len(lab)), b"\0"))
        for t in tracks:
            f.write(t.tobytes())


def load_packed(path):
    """Returns (frames: float32 memmap (total_rows, dim), offsets: int64 (n_tracks + 1), labels: list[str])."""
    with open(path, "rb") as f:
        head = f.read(64)
        if len(head) < 64 or head[:8] != _PK_MAGIC:
            raise IOError("%s is not a packed acoss feature store" % path)
        n, dim, total, lab_bytes = (int(x) for x in np.frombuffer(head[8:40], np.int64))
        if n <= 0 or dim <= 0 or total < 0 or lab_bytes < 0:
            raise IOError("%s: corrupt packed store header" % path)
        offs = np.frombuffer(f.read(8 * (n + 1)), np.int64).copy()
        if len(offs) != n + 1 or offs[0] != 0 or offs[-1] != total or np.any(np.diff(offs) < 0):
            raise IOError("%s: corrupt packed store offsets" % path)
        pos = _pad64(64 + 8 * (n + 1))
        f.seek(pos)
        labels = f.read(lab_bytes).decode("utf-8").split("\n")
        if len(labels) != n:
            raise IOError("%s: corrupt packed store labels" % path)
        data_off = pos + _pad64(lab_bytes)
    frames = np.memmap(path, dtype=np.float32, mode="r", offset=data_off, shape=(total, dim))
    return frames, offs, labels
