# -*- coding: utf-8 -*-
"""
CPU ORACLE -- TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this module.  The product package
(``acoss-1_b200/``) never imports anything under ``oracle/``.

ctypes front-end of ``oracle/acoss_oracle.c`` (the C restatement of essentia's
ChromaCrossSimilarity / CoverSongSimilarity under SURVEY.md Appendix A, *parity
unpinned*, and of the reference's constrained Smith-Waterman, pinned against the
reference's numba code) plus numpy restatements of the small host-side pieces of
the path:

* ``sync_median``          librosa.util.sync(aggregate=np.median), librosa 0.6.1
                           (call sites acoss/algorithms/rqa_serra09.py:51,
                           latefusion_chen.py:51-54)
* ``serra09_normalize``    acoss/algorithms/rqa_serra09.py:71-83
* ``chen_normalize``       acoss/algorithms/latefusion_chen.py:75-85
* ``eval_statistics``      acoss/algorithms/algorithm_template.py:206-291
"""
import ctypes
import os
import subprocess
from concurrent.futures import ThreadPoolExecutor

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build(force=False):
    """Compile oracle/liboracle.so with gcc (oracle/Makefile)."""
    so = os.path.join(_HERE, "liboracle.so")
    src = os.path.join(_HERE, "acoss_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", _HERE, "liboracle.so"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        so = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(so):
            build()
        L = ctypes.CDLL(so)
        f32p = ctypes.POINTER(ctypes.c_float)
        u8p = ctypes.POINTER(ctypes.c_uint8)
        i32p = ctypes.POINTER(ctypes.c_int32)
        i64p = ctypes.POINTER(ctypes.c_int64)
        L.oracle_oti.restype = ctypes.c_int
        L.oracle_oti.argtypes = [f32p, ctypes.c_int, f32p, ctypes.c_int]
        L.oracle_embedded_len.restype = ctypes.c_int
        L.oracle_embedded_len.argtypes = [ctypes.c_int] * 3
        L.oracle_percentile.restype = ctypes.c_float
        L.oracle_percentile.argtypes = [f32p, ctypes.c_int, ctypes.c_float]
        L.oracle_crp.restype = ctypes.c_int
        L.oracle_crp.argtypes = [f32p, ctypes.c_int, f32p, ctypes.c_int,
                                 ctypes.c_int, ctypes.c_int, ctypes.c_float, ctypes.c_int,
                                 u8p, f32p, f32p, f32p, i32p, i64p]
        L.oracle_cover_similarity.restype = ctypes.c_float
        L.oracle_cover_similarity.argtypes = [u8p, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                              ctypes.c_float, ctypes.c_float, f32p]
        L.oracle_sw_constrained.restype = ctypes.c_double
        L.oracle_sw_constrained.argtypes = [u8p, ctypes.c_int, ctypes.c_int]
        L.oracle_serra09_pairs.restype = ctypes.c_int
        L.oracle_serra09_pairs.argtypes = [f32p, i64p, i32p, ctypes.c_int64,
                                           ctypes.c_int, ctypes.c_int, ctypes.c_float, ctypes.c_int,
                                           ctypes.c_float, ctypes.c_float, f32p, f32p, i64p]
        _LIB = L
    return _LIB


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _p(a, ct):
    return a.ctypes.data_as(ctypes.POINTER(ct))


# ----------------------------------------------------------------------------
# essentia restatement
# ----------------------------------------------------------------------------

def oti(query, reference):
    q, r = _f32(query), _f32(reference)
    return int(lib().oracle_oti(_p(q, ctypes.c_float), q.shape[0], _p(r, ctypes.c_float), r.shape[0]))


def embedded_len(n, m=9, tau=1):
    return int(lib().oracle_embedded_len(int(n), int(m), int(tau)))


def percentile(v, kappa):
    v = _f32(v)
    return float(lib().oracle_percentile(_p(v, ctypes.c_float), v.size, ctypes.c_float(kappa)))


def chroma_cross_similarity(query, reference, m=9, tau=1, kappa=0.095, oti=True, full=False):
    """ChromaCrossSimilarity(frameStackSize=m, frameStackStride=tau,
    binarizePercentile=kappa, oti=oti)(query, reference) -> (M, N) uint8 CRP.
    With full=True also returns dict(dist, thr_q, thr_r, oti, n_clamped)."""
    q, r = _f32(query), _f32(reference)
    assert q.ndim == 2 and q.shape[1] == 12 and r.ndim == 2 and r.shape[1] == 12
    M, N = embedded_len(q.shape[0], m, tau), embedded_len(r.shape[0], m, tau)
    if M <= 0 or N <= 0:
        raise RuntimeError("ChromaCrossSimilarity: track too short to stack")
    crp = np.empty((M, N), np.uint8)
    dist = np.empty((M, N), np.float32) if full else None
    tq = np.empty(M, np.float32) if full else None
    tr = np.empty(N, np.float32) if full else None
    o = ctypes.c_int32(0)
    nc = ctypes.c_int64(0)
    rc = lib().oracle_crp(_p(q, ctypes.c_float), q.shape[0], _p(r, ctypes.c_float), r.shape[0],
                          m, tau, ctypes.c_float(kappa), int(bool(oti)),
                          _p(crp, ctypes.c_uint8),
                          _p(dist, ctypes.c_float) if full else None,
                          _p(tq, ctypes.c_float) if full else None,
                          _p(tr, ctypes.c_float) if full else None,
                          ctypes.byref(o), ctypes.byref(nc))
    if rc != 0:
        raise RuntimeError("oracle_crp failed")
    if full:
        return crp, dict(dist=dist, thr_q=tq, thr_r=tr, oti=o.value, n_clamped=nc.value)
    return crp


def cover_song_similarity(crp, alignment="serra09", dis_onset=0.5, dis_extension=0.5, want_matrix=False):
    """CoverSongSimilarity(alignmentType, distanceType='symmetric')(csm) -> (scoreMatrix|None, distance)."""
    R = np.ascontiguousarray(crp, dtype=np.uint8)
    M, N = R.shape
    S = np.empty((M, N), np.float32) if want_matrix else None
    a = {"serra09": 0, "chen17": 1}[alignment]
    v = lib().oracle_cover_similarity(_p(R, ctypes.c_uint8), M, N, a,
                                      ctypes.c_float(dis_onset), ctypes.c_float(dis_extension),
                                      _p(S, ctypes.c_float) if want_matrix else None)
    return S, float(v)


def sw_constrained(B):
    B = np.ascontiguousarray(B, dtype=np.uint8)
    return float(lib().oracle_sw_constrained(_p(B, ctypes.c_uint8), B.shape[0], B.shape[1]))


def serra09_pairs(tracks, pairs, m=9, tau=1, kappa=0.095, oti=True, dis_onset=0.5, dis_extension=0.5,
                  want_dmax=False, n_threads=1):
    """Qmax (and Dmax) raw maxima for a list of pairs.  tracks: list of (n_t,12) f32 arrays.
    n_threads > 1 fans out over a thread pool (ctypes releases the GIL); each call
    is single-threaded like essentia.  Returns (qmax, dmax|None, cells)."""
    frames = _f32(np.concatenate([_f32(t) for t in tracks], axis=0))
    offs = np.zeros(len(tracks) + 1, np.int64)
    offs[1:] = np.cumsum([t.shape[0] for t in tracks])
    pairs = np.ascontiguousarray(pairs, dtype=np.int32).reshape(-1, 2)
    P = pairs.shape[0]
    out_q = np.zeros(P, np.float32)
    out_d = np.zeros(P, np.float32) if want_dmax else None

    def run(lo, hi):
        if hi <= lo:
            return 0
        cells = ctypes.c_int64(0)
        sub = np.ascontiguousarray(pairs[lo:hi])
        oq = out_q[lo:hi]
        od = out_d[lo:hi] if want_dmax else None
        rc = lib().oracle_serra09_pairs(_p(frames, ctypes.c_float), _p(offs, ctypes.c_int64),
                                        _p(sub, ctypes.c_int32), hi - lo, m, tau, ctypes.c_float(kappa),
                                        int(bool(oti)), ctypes.c_float(dis_onset), ctypes.c_float(dis_extension),
                                        _p(oq, ctypes.c_float), _p(od, ctypes.c_float) if want_dmax else None,
                                        ctypes.byref(cells))
        if rc != 0:
            raise RuntimeError("oracle_serra09_pairs: pair %d has a too-short track" % (lo - rc - 1))
        return cells.value

    if n_threads <= 1 or P < 2:
        cells = run(0, P)
    else:
        # interleave-free contiguous chunks, a few per thread for balance
        nchunk = min(P, n_threads * 4)
        bounds = np.linspace(0, P, nchunk + 1).astype(np.int64)
        with ThreadPoolExecutor(n_threads) as ex:
            cells = sum(ex.map(lambda ab: run(int(ab[0]), int(ab[1])), zip(bounds[:-1], bounds[1:])))
    return out_q, out_d, cells


# ----------------------------------------------------------------------------
# host-side pieces of the path, restated in numpy
# ----------------------------------------------------------------------------

def sync_median(chroma, fac):
    """librosa.util.sync(chroma.T, np.arange(0, T, fac), aggregate=np.median).T
    (librosa 0.6.1, pad=True): boundaries unique({0} U idx U {T}); one output frame
    per segment [b_k, b_{k+1}); the last one may be short.  chroma: (T, 12)."""
    x = np.asarray(chroma)
    T = x.shape[0]
    b = np.unique(np.concatenate([[0], np.arange(0, T, fac), [T]])).astype(int)
    out = np.empty((len(b) - 1, x.shape[1]), dtype=x.dtype)
    for k in range(len(b) - 1):
        out[k] = np.median(x[b[k]:b[k + 1]], axis=0)
    return out


def serra09_normalize(D, lengths):
    """Ds[i, j] /= sqrt(n_j), float32 store (rqa_serra09.py:71-83)."""
    D = np.array(D, dtype=np.float32)
    for j in range(D.shape[1]):
        nf = np.sqrt(lengths[j])
        for i in range(D.shape[0]):
            D[i, j] /= nf
    return D


def chen_normalize(D, lengths):
    """Ds[i, j] = sqrt(n_j) / Ds[i, j], float32 store (latefusion_chen.py:75-85)."""
    D = np.array(D, dtype=np.float32)
    with np.errstate(divide="ignore"):
        for j in range(D.shape[1]):
            nf = np.sqrt(lengths[j])          # float64 scalar; f64 / f32 -> f64, stored as f32
            for i in range(D.shape[0]):
                D[i, j] = nf / D[i, j]
    return D


def eval_statistics(D, cliques, topsidx=(1, 10, 100, 1000), stable=True):
    """getEvalStatistics restated (algorithm_template.py:206-291), without the
    prints / CSV.  cliques: dict label -> iterable of indices, in the reference's
    insertion order.  stable=True ranks ties by clique-reordered index (the
    reference's np.argsort default is unstable: SURVEY Q10/H5)."""
    from itertools import chain
    D = np.array(D, dtype=np.float32)
    N = D.shape[0]
    cl = [list(cliques[s]) for s in cliques]
    Ks = np.array([len(c) for c in cl])
    idx = np.argsort(-Ks, kind="stable" if stable else None)
    Ks = Ks[idx]
    cl = [cl[i] for i in idx]
    idx = np.array(list(chain(*cl)), dtype=int)
    D = D[idx, :][:, idx]
    np.fill_diagonal(D, -np.inf)
    order = np.argsort(-D, 1, kind="stable" if stable else None)
    ranks = np.nan * np.ones(N)
    AllMap = np.nan * np.ones(N)
    startidx = 0
    kidx = 0
    for i in range(N):
        if i >= startidx + Ks[kidx]:
            startidx += Ks[kidx]
            kidx += 1
            if Ks[kidx] < 2:
                break
        diff = order[i] - startidx
        iranks = (np.nonzero((diff >= 0) & (diff < Ks[kidx]))[0] + 1).tolist()
        iranks = iranks[0:-1]
        if len(iranks) == 0:
            break
        ranks[i] = iranks[0]
        P = np.array([float(j) / float(r) for (j, r) in zip(range(1, Ks[kidx]), iranks)])
        AllMap[i] = np.mean(P)
    MAP = np.nanmean(AllMap)
    ranks = ranks[np.isnan(ranks) == 0]
    MR = np.mean(ranks)
    MRR = 1.0 / N * (np.sum(1.0 / ranks))
    MDR = np.median(ranks)
    tops = np.array([np.sum(ranks <= t) for t in topsidx], dtype=np.float64)
    return MR, MRR, MDR, MAP, tops
