# -*- coding: utf-8 -*-
"""
CPU ORACLE (reference loader) -- TEST INFRASTRUCTURE ONLY, THIS CONTAINER ONLY.

Imports the *unmodified* reference from /root/reference read-only, with stub
modules for the third-party packages that are absent from this image
(essentia, librosa, deepdish, progress, skimage).  /root/reference does not
exist on the GPU box, so nothing at GPU run time may import this file: it is
used by ``oracle/make_golden.py`` to freeze reference outputs under
``tests/golden/`` and by the ``-m "not gpu"`` tests when the reference is present
(they skip otherwise).

What runs verbatim through this loader:
  CoverAlgorithm (incl. getEvalStatistics), smith_waterman_constrained (numba
  compiled), cross_recurrence.* (with numba.jit := identity, because its
  nopython decorators do not type-check), similarity_fusion.*, Simple, FTM2D,
  and the Serra09 / ChenFusion *orchestration* on top of the C restatement of
  essentia (oracle.py), plugged in as the stub ``essentia.standard``.
"""
import os
import pickle
import sys
import types

import numpy as np

REF = "/root/reference"


def available():
    return os.path.isdir(os.path.join(REF, "acoss"))


class _DeepdishIO(object):
    """deepdish.io stand-in with the same load/save API.  Reads the npz container the synthetic generator writes
    (acoss_b200.featio: '/'-joined keys, no pickle); files the reference classes write THEMSELVES through this stub
    (their private caches in the test's temporary directory) are pickles."""

    @staticmethod
    def load(path):
        with open(path, "rb") as f:
            head = f.read(4)
        if head == b"PK\x03\x04":
            root = {}
            with np.load(path, allow_pickle=False) as z:
                for key in z.files:
                    if key.startswith("__"):
                        continue
                    parts = key.split("/")
                    d = root
                    for q in parts[:-1]:
                        d = d.setdefault(q, {})
                    if parts[-1] != "__empty__":
                        a = z[key]
                        d[parts[-1]] = a.item() if a.ndim == 0 else a
            return root
        with open(path, "rb") as f:
            return pickle.load(f)

    @staticmethod
    def save(path, obj):
        if isinstance(obj, dict):
            obj = {k: np.array(v) if isinstance(v, np.memmap) else v for k, v in obj.items()}
        with open(path, "wb") as f:
            pickle.dump(obj, f, protocol=4)


class _Bar(object):
    def __init__(self, *a, **k):
        pass

    def next(self):
        pass

    def finish(self):
        pass


def _install_stubs(identity_jit_for_cross_recurrence=True):
    from . import oracle as O

    # --- essentia.standard backed by the C restatement -----------------------
    ess = types.ModuleType("essentia")
    std = types.ModuleType("essentia.standard")

    class ChromaCrossSimilarity(object):
        def __init__(self, frameStackSize=9, frameStackStride=1, binarizePercentile=0.095, oti=True, **kw):
            self.m, self.tau, self.kappa, self.oti = frameStackSize, frameStackStride, binarizePercentile, oti

        def __call__(self, query, reference):
            return O.chroma_cross_similarity(query, reference, self.m, self.tau, self.kappa, self.oti).astype(np.float32)

    class CoverSongSimilarity(object):
        def __init__(self, alignmentType="serra09", distanceType="asymmetric", disOnset=0.5, disExtension=0.5, **kw):
            self.a, self.go, self.ge = alignmentType, disOnset, disExtension

        def __call__(self, csm):
            return O.cover_song_similarity(np.asarray(csm).astype(np.uint8), self.a, self.go, self.ge, want_matrix=False)

    std.ChromaCrossSimilarity = ChromaCrossSimilarity
    std.CoverSongSimilarity = CoverSongSimilarity
    ess.standard = std
    sys.modules.setdefault("essentia", ess)
    sys.modules.setdefault("essentia.standard", std)

    # --- librosa -------------------------------------------------------------
    lib = types.ModuleType("librosa")
    util = types.ModuleType("librosa.util")
    filt = types.ModuleType("librosa.filters")

    def sync(data, idx, aggregate=None, pad=True, axis=-1):
        assert axis == -1 and pad
        agg = aggregate or np.mean
        T = data.shape[-1]
        b = np.unique(np.concatenate([[0], np.asarray(idx, dtype=int), [T]])).astype(int)
        b = b[(b >= 0) & (b <= T)]
        out = np.empty(data.shape[:-1] + (len(b) - 1,), dtype=data.dtype)
        for k in range(len(b) - 1):
            out[..., k] = agg(data[..., b[k]:b[k + 1]], axis=-1)
        return out

    def normalize(S, norm=np.inf, axis=0, threshold=None, fill=None):
        assert norm == 2
        mag = np.abs(S).astype(float)
        length = np.sum(mag ** 2, axis=axis, keepdims=True) ** 0.5
        tiny = np.finfo(S.dtype if np.issubdtype(S.dtype, np.floating) else np.float32).tiny
        small = length < tiny
        length = np.where(small, 1.0, length)
        return S / length

    def get_window(window, Nx, fftbins=True):
        import scipy.signal
        return scipy.signal.get_window(window, Nx, fftbins=fftbins)

    util.sync = sync
    util.normalize = normalize
    filt.get_window = get_window
    lib.util = util
    lib.filters = filt
    sys.modules.setdefault("librosa", lib)
    sys.modules.setdefault("librosa.util", util)
    sys.modules.setdefault("librosa.filters", filt)

    # --- deepdish, progress ---------------------------------------------------
    dd = types.ModuleType("deepdish")
    dd.io = _DeepdishIO
    sys.modules.setdefault("deepdish", dd)
    sys.modules.setdefault("deepdish.io", dd.io)
    prog = types.ModuleType("progress")
    bar = types.ModuleType("progress.bar")
    bar.Bar = _Bar
    prog.bar = bar
    sys.modules.setdefault("progress", prog)
    sys.modules.setdefault("progress.bar", bar)


_LOADED = {}


def load():
    """Returns a namespace with the reference's modules:
    .template (algorithm_template), .align (alignment_tools), .xrec (cross_recurrence,
    identity-jit), .snf (similarity_fusion), .serra09, .chen, .simple, .ftm2d, .early"""
    if _LOADED:
        return _LOADED["ns"]
    if not available():
        raise RuntimeError("/root/reference is not present")
    _install_stubs()
    import importlib
    import importlib.util

    # Build a synthetic top-level package 'acoss' WITHOUT running acoss/__init__.py
    # (it imports features.py / extractors.py = audio front-end with many absent deps).
    pkg = types.ModuleType("acoss")
    pkg.__path__ = [os.path.join(REF, "acoss")]
    sys.modules["acoss"] = pkg
    cwd = os.getcwd()
    try:
        # utils.log() opens a log file in cwd: keep that out of the repo
        import tempfile
        os.chdir(tempfile.gettempdir())
        utils = importlib.import_module("acoss.utils")
        algs = types.ModuleType("acoss.algorithms")
        algs.__path__ = [os.path.join(REF, "acoss", "algorithms")]
        sys.modules["acoss.algorithms"] = algs
        autils = types.ModuleType("acoss.algorithms.utils")
        autils.__path__ = [os.path.join(REF, "acoss", "algorithms", "utils")]
        sys.modules["acoss.algorithms.utils"] = autils

        align = importlib.import_module("acoss.algorithms.utils.alignment_tools")   # numba, compiles
        # cross_recurrence: numba.jit -> identity (its nopython decorators fail typing)
        import numba
        real_jit = numba.jit
        numba.jit = lambda *a, **k: (a[0] if (len(a) == 1 and callable(a[0]) and not k) else (lambda f: f))
        try:
            xrec = importlib.import_module("acoss.algorithms.utils.cross_recurrence")
        finally:
            numba.jit = real_jit
        snf = importlib.import_module("acoss.algorithms.utils.similarity_fusion")
        template = importlib.import_module("acoss.algorithms.algorithm_template")
        serra = importlib.import_module("acoss.algorithms.rqa_serra09")
        chen = importlib.import_module("acoss.algorithms.latefusion_chen")
        simple = importlib.import_module("acoss.algorithms.simple_silva")
        ftm = importlib.import_module("acoss.algorithms.ftm2d")
        # Q5: ftm2d.py uses os, dd, scipy.fftpack without importing them
        import scipy.fftpack
        ftm.os = os
        ftm.dd = sys.modules["deepdish"]
        ftm.scipy = __import__("scipy")
        early = importlib.import_module("acoss.algorithms.earlyfusion_traile")
        early.os = os
        import time as _time
        early.time = _time
    finally:
        os.chdir(cwd)
    ns = types.SimpleNamespace(utils=utils, template=template, align=align, xrec=xrec, snf=snf,
                               serra09=serra, chen=chen, simple=simple, ftm2d=ftm, early=early,
                               dd=sys.modules["deepdish"])
    _LOADED["ns"] = ns
    return ns
