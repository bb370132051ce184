# -*- coding: utf-8 -*-
"""
Per-track feature files and result files.  The reference stores deepdish-HDF5 dicts
(`dd.io.load(path)` / `dd.io.save`, acoss/algorithms/algorithm_template.py:90,193;
schema README.md:94-114).

* With deepdish importable, `save` writes real deepdish HDF5 and `load` reads it: the files
  are then interchangeable with the reference's.
* deepdish / h5py / pytables are NOT installed in this image, so here `save` writes the same
  (nested) dict under the same `.h5` path as a numpy `.npz` archive with '/'-joined keys --
  a plain data container, loaded with `allow_pickle=False`.  A real Da-TACOS `.h5` directory
  therefore cannot be read on this box; `load` says so instead of guessing.
* Nothing is ever unpickled: a feature directory is untrusted input.
"""
import numpy as np

try:  # pragma: no cover - not available in this image
    import deepdish as _dd
    if not hasattr(_dd, "io") or getattr(_dd.io, "__name__", "") == "_DeepdishIO":
        _dd = None
except Exception:  # ImportError and friends
    _dd = None

_HDF5_MAGIC = b"\x89HDF\r\n\x1a\n"
_ZIP_MAGIC = b"PK\x03\x04"
_NPZ_MARK = "__acoss_npz__"


def _flatten(obj, prefix, out):
    if isinstance(obj, dict):
        if not obj:
            out[prefix + "/__empty__" if prefix else "__empty__"] = np.zeros(0, np.uint8)
        for k, v in obj.items():
            k = str(k)
            if "/" in k or k.startswith("__"):
                raise ValueError("featio.save: key %r cannot be stored" % k)
            _flatten(v, prefix + "/" + k if prefix else k, out)
        return
    a = np.array(obj) if isinstance(obj, np.memmap) else np.asarray(obj)
    if a.dtype == object:
        raise ValueError("featio.save: value under %r is not plain numeric / string data" % prefix)
    out[prefix] = a


def _unflatten(npz):
    root = {}
    for key in npz.files:
        if key == _NPZ_MARK:
            continue
        parts = key.split("/")
        d = root
        for p in parts[:-1]:
            d = d.setdefault(p, {})
        if parts[-1] == "__empty__":
            continue
        a = npz[key]
        if a.ndim == 0:
            a = a.item()            # str / int / float leaves come back as Python scalars, like deepdish
        d[parts[-1]] = a
    return root


def load(path):
    with open(path, "rb") as f:
        head = f.read(8)
    if head == _HDF5_MAGIC:
        if _dd is None:
            raise IOError("%s is an HDF5 file but deepdish is not installed in this environment" % path)
        return _dd.io.load(path)
    if head[:4] == _ZIP_MAGIC:
        with np.load(path, allow_pickle=False) as z:
            if _NPZ_MARK not in z.files:
                raise IOError("%s is a zip archive but not an acoss_b200 feature file" % path)
            return _unflatten(z)
    raise IOError("%s is neither deepdish HDF5 nor an acoss_b200 npz container (pickled files are refused: "
                  "feature directories are untrusted input)" % path)


def save(path, obj):
    if isinstance(obj, dict):
        obj = {k: (np.array(v) if isinstance(v, np.memmap) else v) for k, v in obj.items()}
    if _dd is not None:  # pragma: no cover - not available in this image
        _dd.io.save(path, obj)
        return
    if not isinstance(obj, dict):
        raise ValueError("featio.save: only dicts are stored")
    flat = {}
    _flatten(obj, "", flat)
    flat[_NPZ_MARK] = np.array(1, np.uint8)
    with open(path, "wb") as f:      # a file object keeps the `.h5` name (np.savez would append .npz to a str)
        np.savez(f, **flat)


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
