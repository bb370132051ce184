# -*- coding: utf-8 -*-
"""
Similarity network fusion (Wang et al. 2012/2014): host-side mirror of
acoss/algorithms/utils/similarity_fusion.py `doSimilarityFusion` (:188-196).  All arithmetic
(getW, getP, getS, the cross-diffusion iterations with the reference's in-place aliasing) runs on
the GPU through the C-ABI entry point `acoss_snf`; there is no CPU path.
"""
import numpy as np

__all__ = ['doSimilarityFusion']


def doSimilarityFusion(Scores, K=5, niters=5, reg_diag=1, ctx=None, device=0):
    """Do similarity fusion on a set of NxN distance matrices.
    :returns (An array of similarity matrices for each feature, Fused Similarity Matrix)"""
    if ctx is None:
        from ._lib import Context
        ctx = Context(device)
    Ws, fused = ctx.snf([np.asarray(D) for D in Scores], K=K, niters=niters, reg_diag=reg_diag)
    return (Ws, fused)
