# -*- coding: utf-8 -*-
"""
Tralie, C.J., 2017. Early MFCC and HPCP fusion for robust cover song identification (ISMIR 2017).

Host-side mirror of the reference's `EarlyFusion` (acoss/algorithms/earlyfusion_traile.py:44-206):
same constructor keywords, similarity types ("mfccs", "ssms", "chromas", "early"), cache prefix,
`load_features` (in-memory / on-disk cache of the blocked features), `similarity(idxs)` and
`do_late_fusion()`.  The whole per-pair body (3 cross-similarity GEMMs, kNN binarisation, constrained
Smith-Waterman x4, kernel early fusion) runs on the GPU through acoss_earlyfusion_pairs; the N x N
similarity network fusion through acoss_snf.

The blocked features (`load_features`, earlyfusion_traile.py:104-150: beat-synchronous blocks, anti-aliased
resize, z-normalisation, per-block SSM, chroma blocks and medians) are computed on the GPU too
(acoss_earlyfusion_block_features), for all uncached tracks in one batch (`preload_features`), and written to
the reference's cache files `<prefix>_<i>.h5`.  `resize_block` is skimage.transform.resize restated (skimage is
absent here: that stage's parity is unpinned).
"""
import os

import numpy as np

from . import featio
from .algorithm_template import CoverAlgorithm
from .similarity_fusion import doSimilarityFusion

__all__ = ['EarlyFusion']


class EarlyFusion(CoverAlgorithm):
    def __init__(self, dataset_csv, datapath, chroma_type='hpcp', shortname='Covers80', blocksize=20,
                 mfccs_per_block=50, ssm_res=50, chromas_per_block=40, kappa=0.1, K=10, niters=5, log_times=False,
                 cachedir="cache", device=None):
        self.chroma_type = chroma_type
        self.blocksize = blocksize
        self.mfccs_per_block = mfccs_per_block
        self.chromas_per_block = chromas_per_block
        self.kappa = kappa
        self.K = K
        self.niters = niters
        self.all_block_feats = {}  # Cached features
        self.log_times = log_times
        if log_times:
            self.times = {'features': [], 'raw': []}
        self._uploaded = False
        CoverAlgorithm.__init__(self, dataset_csv=dataset_csv, name="EarlyFusionTraile", datapath=datapath,
                                shortname=shortname, similarity_types=["mfccs", "ssms", "chromas", "early"],
                                cachedir=cachedir, device=device)

    def get_cacheprefix(self):
        return "%s/%s_%s_%s" % (self.cachedir, self.name, self.shortname, self.chroma_type)

    def set_block_features(self, i, block_feats):
        """Inject precomputed blocked features of song i (and write the reference's cache file)."""
        self.all_block_feats[i] = block_feats
        featio.save("%s_%i.h5" % (self.get_cacheprefix(), i), block_feats)
        self._uploaded = False

    def preload_features(self, indices=None):
        """Blocked features of every track that is neither in memory nor in the on-disk cache, computed on the device in one
        batch and written to the reference's cache files."""
        todo = []
        for i in (range(self.N) if indices is None else indices):
            if i in self.all_block_feats:
                continue
            filepath = "%s_%i.h5" % (self.get_cacheprefix(), i)
            if os.path.exists(filepath):
                self.all_block_feats[i] = featio.load(filepath)
                CoverAlgorithm.load_features(self, i)      # clique info as a side effect
            else:
                todo.append(i)
        if not todo:
            return
        mf, ch, on = [], [], []
        for i in todo:
            feats = CoverAlgorithm.load_features(self, i)
            mf.append(np.ascontiguousarray(np.asarray(feats['mfcc_htk'], np.float32).T))
            ch.append(np.asarray(feats[self.chroma_type], np.float32))
            on.append(np.asarray(feats['madmom_features']['onsets']))
        out = self.context().earlyfusion_block_features(mf, ch, on, blocksize=self.blocksize, mfccs_per_block=self.mfccs_per_block,
                                                        chromas_per_block=self.chromas_per_block)
        for i, bf in zip(todo, out):
            self.all_block_feats[i] = bf
            featio.save("%s_%i.h5" % (self.get_cacheprefix(), i), bf)
        self._uploaded = False

    def load_features(self, i, do_plot=False):
        if i not in self.all_block_feats:
            self.preload_features([i])
        return self.all_block_feats[i]

    def preload(self):
        if self._uploaded:
            return
        self.preload_features()
        self.context().earlyfusion_upload([self.load_features(i) for i in range(self.N)])
        self._uploaded = True

    def pair_cost(self, idxs):
        n = np.array([self.load_features(i)['mfccs'].shape[0] for i in range(self.N)], dtype=np.int64)
        return n[idxs[:, 0]] * n[idxs[:, 1]]

    def similarity(self, idxs, do_plot=False):
        idxs = np.asarray(idxs).reshape(-1, 2)
        if len(idxs) == 0:
            return
        self.preload()
        ctx = self.context()
        scores = ctx.earlyfusion_pairs(idxs, kappa=self.kappa, K=self.K)
        self.last_stats = ctx.earlyfusion_stats()
        for k, s in enumerate(["mfccs", "ssms", "chromas", "early"]):
            self.Ds[s][idxs[:, 0], idxs[:, 1]] = scores[:, k]

    def do_late_fusion(self):
        """Late fusion after all pairwise scores have been computed (earlyfusion_traile.py:200-206)."""
        ctx, dev = self.context(), self.device_index()
        self.Ds["late"] = doSimilarityFusion([1.0 / (1.0 + self.Ds[s]) for s in ["chromas", "ssms", "mfccs"]],
                                             K=20, niters=20, reg_diag=1, ctx=ctx, device=dev)[1]
        self.Ds["early+late"] = doSimilarityFusion([1.0 / (1.0 + self.Ds[s]) for s in ["chromas", "ssms", "mfccs", "early"]],
                                                   K=20, niters=20, reg_diag=1, ctx=ctx, device=dev)[1]
