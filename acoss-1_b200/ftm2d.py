# -*- coding: utf-8 -*-
"""
Ellis, D. P., & Bertin-Mahieux, T. (2012). Large-scale cover song recognition using the 2D Fourier
transform magnitude (ISMIR 2012).

Host-side mirror of the reference's `FTM2D` (acoss/algorithms/ftm2d.py:23-97): same constructor
keywords, cache prefix, `load_features(i)` (returns the 900-d unit shingle, cached in memory and in
`<prefix>_<i>.h5` like the reference) and `similarity(idxs)`.

Both stages run on the GPU through the C-ABI: the per-track shingles (beat-synchronous median chroma ->
chrompwr -> 12 x WIN 2-D DFT magnitude patches -> log compression -> median -> unit norm, ftm2d.py:52-66,
100-138) for ALL tracks of the dataset in one batch (`preload`, acoss_ftm2d_shingles), and the pair scores
exp(-||s_i - s_j||^2) (:85-97) from one all-pairs kernel (acoss_ftm2d_allpairs).  There is no host
arithmetic left in this module; the numpy restatement of the reference lives in oracle/ports.py (tests).
"""
import os

import numpy as np

from . import featio
from .algorithm_template import CoverAlgorithm

__all__ = ['FTM2D']


class FTM2D(CoverAlgorithm):
    """
    Attributes
    ----------
    Same as CoverAlgorithms, plus
    shingles: {int: ndarray(WIN*chromabins)}
        A map from the song index to the FFT2DM shingles, so that they are cached
    chroma_type: string
        Type of chroma to use (key into features)
    """

    def __init__(self, dataset_csv, datapath, chroma_type='hpcp', shortname='Covers80', PWR=1.96, WIN=75, C=5,
                 cachedir="cache", device=None):
        self.PWR = PWR
        self.WIN = WIN
        self.C = C
        self.chroma_type = chroma_type
        self.shingles = {}
        self._D = None
        CoverAlgorithm.__init__(self, dataset_csv=dataset_csv, name="FTM2D", datapath=datapath,
                                shortname=shortname, cachedir=cachedir, device=device)

    def get_cacheprefix(self):
        return "%s/%s_%s_%s" % (self.cachedir, self.name, self.shortname, self.chroma_type)

    def preload(self, indices=None):
        """Shingles of every track that is neither in memory nor in the on-disk cache, computed on the device in one
        batch and written to the reference's cache files."""
        todo = []
        for i in (range(self.N) if indices is None else indices):
            if i in self.shingles:
                continue
            filepath = "%s_%i.h5" % (self.get_cacheprefix(), i)
            if os.path.exists(filepath):
                self.shingles[i] = np.asarray(featio.load(filepath)['shingle'], np.float64)
                CoverAlgorithm.load_features(self, i)      # clique info as a side effect
            else:
                todo.append(i)
        if not todo:
            return
        raw, onsets = [], []
        for i in todo:
            feats = CoverAlgorithm.load_features(self, i)
            raw.append(np.asarray(feats[self.chroma_type], np.float32))
            onsets.append(np.asarray(feats['madmom_features']['onsets']))
        S = self.context().ftm2d_shingles(raw, onsets, pwr=self.PWR, win=self.WIN, C=self.C)
        for k, i in enumerate(todo):
            self.shingles[i] = S[k]
            featio.save("%s_%i.h5" % (self.get_cacheprefix(), i), {'shingle': S[k]})

    def load_features(self, i, do_plot=False):
        if i not in self.shingles:
            self.preload([i])
        return self.shingles[i]

    def similarity(self, idxs):
        idxs = np.asarray(idxs).reshape(-1, 2)
        if len(idxs) == 0:
            return
        if self._D is None:
            self.preload()
            S = np.stack([np.asarray(self.shingles[i], dtype=np.float64) for i in range(self.N)])
            self._D = self.context().ftm2d_allpairs(S)
        self.Ds['main'][idxs[:, 0], idxs[:, 1]] = self._D[idxs[:, 0], idxs[:, 1]]
