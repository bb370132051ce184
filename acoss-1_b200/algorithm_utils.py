# -*- coding: utf-8 -*-
"""
The free functions of the reference's `acoss.algorithms.utils` package (alignment_tools.py, cross_recurrence.py,
similarity_fusion.py) under their own names, so that a third-party CoverAlgorithm subclass that calls them keeps working.
Every function hands its matrices to the C-ABI library (csrc/libacoss_b200.so) and returns what the device computed; there is no
host arithmetic here and no CPU path -- without the library (or a GPU) the call raises.

All functions take an optional `ctx` (an `acoss_b200.Context`); without one a process-wide context on device 0 is created on first
use (`set_default_context` replaces it).
"""
import numpy as np

__all__ = ["smith_waterman_constrained", "get_ssm", "get_csm", "get_csm_cosine", "get_oti", "get_csm_blocked_oti", "csm_to_binary",
           "getW", "getWCSM", "getP", "getS", "doSimilarityFusionWs", "doSimilarityFusion", "set_default_context"]

_default = None


def set_default_context(ctx):
    """The context the free functions use when none is passed."""
    global _default
    _default = ctx


def _ctx(ctx):
    global _default
    if ctx is not None:
        return ctx
    if _default is None:
        from ._lib import Context
        _default = Context(0)
    return _default


# ---- alignment_tools.py -----------------------------------------------------------------------------------------------
def smith_waterman_constrained(input_matrix, ctx=None):
    """Constrained Smith-Waterman score of one binary matrix (alignment_tools.py:25-46)."""
    B = np.ascontiguousarray(np.asarray(input_matrix) != 0, dtype=np.uint8)
    return float(_ctx(ctx).sw_constrained([B])[0])


# ---- cross_recurrence.py ----------------------------------------------------------------------------------------------
def get_csm(X, Y, ctx=None):
    """Euclidean cross-similarity of two point clouds (M, d), (N, d) -> (M, N) (cross_recurrence.py:31-48)."""
    return _ctx(ctx).csm(X, Y, cosine=False)


def get_ssm(X, ctx=None):
    """Euclidean self-similarity of a point cloud (cross_recurrence.py:11-28): get_csm(X, X) with the diagonal zeroed."""
    D = _ctx(ctx).csm(X, X, cosine=False)
    D[np.diag_indices(D.shape[0])] = 0
    return D


def get_csm_cosine(X, Y, ctx=None):
    """Cosine distance 1 - <x, y> / (|x| |y|) between the rows (cross_recurrence.py:54-73)."""
    return _ctx(ctx).csm(X, Y, cosine=True)


def get_oti(C1, C2, do_plot=False, ctx=None):
    """Index by which to rotate chroma vector C1 to match C2 best (cross_recurrence.py:76-103).  12 bins."""
    if do_plot:
        raise NotImplementedError("do_plot: there is no display on this side of the interface")
    return int(_ctx(ctx).oti(np.asarray(C1).reshape(1, -1), np.asarray(C2).reshape(1, -1))[0])


def get_csm_blocked_oti(X, Y, C1, C2, csm_fn, ctx=None):
    """csm_fn(X rolled by the global OTI of (C1, C2), Y) (cross_recurrence.py:106-134).  The roll is an index permutation."""
    nb = len(C1)
    per = int(X.shape[1] / nb)
    oti = get_oti(C1, C2, ctx=ctx)
    X1 = np.roll(np.reshape(X, (X.shape[0], per, nb)), oti, axis=2).reshape(X.shape[0], per * nb)
    if csm_fn in (get_csm, get_csm_cosine):
        return csm_fn(X1, Y, ctx=ctx)
    return csm_fn(X1, Y)


def csm_to_binary(D, kappa, ctx=None):
    """Binary cross-similarity matrix of the kappa nearest columns of every row (cross_recurrence.py:137-161)."""
    return _ctx(ctx).csm_to_binary(D, kappa)


# ---- similarity_fusion.py ---------------------------------------------------------------------------------------------
def getW(D, K, Mu=0.5, ctx=None):
    """Affinity matrix of a distance matrix with the K-nearest-neighbour local scaling (similarity_fusion.py:15-36)."""
    if Mu != 0.5:
        raise NotImplementedError("Mu != 0.5: the reference never passes another value and the device kernel fixes it")
    return _ctx(ctx).snf_stages([D], K, want=("Ws",))["Ws"][0]


def getWCSM(CSMAB, k1, k2, Mu=0.5, ctx=None):
    """Cross-affinity of a cross-similarity matrix (similarity_fusion.py:38-54)."""
    if Mu != 0.5:
        raise NotImplementedError("Mu != 0.5: the reference never passes another value and the device kernel fixes it")
    return _ctx(ctx).wcsm(CSMAB, k1, k2)


def getP(W, diagRegularize=False, ctx=None):
    """Row-normalised transition matrix (similarity_fusion.py:101-122, the diagRegularize=False branch every caller uses)."""
    if diagRegularize:
        raise NotImplementedError("diagRegularize=True is not used by any algorithm of the reference")
    W = np.asarray(W)
    P = _ctx(ctx).snf_stages([W], 1, affinities=True, want=("P",))["P"][0]
    return P.astype(W.dtype) if W.dtype == np.float32 else P


def getS(W, K, ctx=None):
    """Sparse K-nearest-neighbour kernel matrix (similarity_fusion.py:124-144), as scipy CSR like the reference."""
    from scipy import sparse
    S = _ctx(ctx).snf_stages([W], K, affinities=True, want=("S",))["S"][0]
    return sparse.csr_matrix(S)


def doSimilarityFusionWs(Ws, K=5, niters=20, reg_diag=1, ctx=None):
    """Cross-diffusion of a list of affinity matrices (similarity_fusion.py:146-186), with the reference's list aliasing."""
    return _ctx(ctx).snf_stages(list(Ws), K, niters=niters, reg_diag=reg_diag, affinities=True, want=("fused",))["fused"]


def doSimilarityFusion(Scores, K=5, niters=5, reg_diag=1, ctx=None):
    """(Ws, fused) of a list of distance matrices (similarity_fusion.py:188-196)."""
    return _ctx(ctx).snf([np.asarray(D) for D in Scores], K=K, niters=niters, reg_diag=reg_diag)
