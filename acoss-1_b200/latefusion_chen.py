# -*- coding: utf-8 -*-
"""
Chen, N., Li, W. and Xiao, H., 2018. Fusing similarity functions for cover song identification.
Multimedia Tools and Applications, 77(2), pp.2629-2652.

Host-side mirror of the reference's `ChenFusion` (acoss/algorithms/latefusion_chen.py:32-91).
"""
import numpy as np

from .algorithm_template import CoverAlgorithm
from .rqa_serra09 import ChromaPairAlgorithm

__all__ = ['ChenFusion']


class ChenFusion(ChromaPairAlgorithm):
    def __init__(self, dataset_csv, datapath, chroma_type='hpcp', shortname='benchmark',
                 oti=True, kappa=0.095, tau=1, m=9, downsample_fac=40, cachedir="cache", device=None):
        self._init_chroma(chroma_type, oti, kappa, tau, m, downsample_fac)
        CoverAlgorithm.__init__(self, dataset_csv, name="LateFusionChen", similarity_types=["qmax", "dmax"],
                                datapath=datapath, shortname=shortname, cachedir=cachedir, device=device)

    def similarity(self, idxs):
        idxs = np.asarray(idxs).reshape(-1, 2)
        if len(idxs) == 0:
            return
        self.preload()
        ctx = self.context()
        qmax, dmax = ctx.serra09_pairs(idxs, self.crp_params(), want_dmax=True)
        self.last_stats = ctx.pair_stats()
        self.Ds["qmax"][idxs[:, 0], idxs[:, 1]] = qmax
        self.Ds["dmax"][idxs[:, 0], idxs[:, 1]] = dmax

    def normalize_by_length(self):
        """Ds[i, j] = sqrt(len_j) / Ds[i, j] (latefusion_chen.py:75-85): scores become distances,
        zero scores / the diagonal become inf."""
        self._normalize(1)

    def do_late_fusion(self):
        """SNF of the two distance matrices, K=20, niters=20, reg_diag=1 (latefusion_chen.py:87-91)."""
        from .similarity_fusion import doSimilarityFusion
        DLate = doSimilarityFusion([self.Ds[s] for s in self.Ds], K=20, niters=20, reg_diag=1,
                                   ctx=self.context(), device=self.device_index())[1]
        for key in self.Ds:
            self.Ds[key] *= -1  # Switch back to larger scores being closer
        self.Ds["Late"] = DLate
