# -*- coding: utf-8 -*-
"""
Ellis, D. P., & Bertin-Mahieux, T. (2012). Large-scale cover song recognition using the 2D Fourier
transform magnitude (ISMIR 2012).

Host-side mirror of the reference's `FTM2D` (acoss/algorithms/ftm2d.py:23-138): same constructor
keywords, cache prefix, `load_features` (beat-synchronous median chroma -> chrompwr -> 12 x WIN
2D-FFT magnitude patches -> log compression -> median shingle, cached per track like the
reference) and `similarity(idxs)`.  The pair scores exp(-||s_i - s_j||^2) (:85-97) come from one
all-pairs kernel on the GPU (acoss_ftm2d_allpairs).
"""
import os

import numpy as np
import scipy.fftpack

from . import featio
from .algorithm_template import CoverAlgorithm

__all__ = ['FTM2D', 'chrompwr', 'btchroma_to_fftmat']


def _sync_median(data, idx):
    """librosa.util.sync(data, idx, aggregate=np.median) (librosa 0.6.1): segments between the unique
    boundaries {0} u idx u {T} along the last axis."""
    T = data.shape[-1]
    b = np.unique(np.concatenate([[0], np.asarray(idx, dtype=int), [T]])).astype(int)
    b = b[(b >= 0) & (b <= T)]
    out = np.empty(data.shape[:-1] + (len(b) - 1,), dtype=data.dtype)
    for k in range(len(b) - 1):
        out[..., k] = np.median(data[..., b[k]:b[k + 1]], axis=-1)
    return out


def chrompwr(X, P=.5):
    """Raise chroma columns to a power, preserving their norm (ftm2d.py:100-117)."""
    nchr, nbts = X.shape
    CMn = np.tile(np.sqrt(np.sum(X * X, axis=0)), (nchr, 1))
    CMn[CMn == 0] = 1
    CMp = np.power(X / CMn, P)
    CMpn = np.tile(np.sqrt(np.sum(CMp * CMp, axis=0)), (nchr, 1))
    CMpn[np.where(CMpn == 0)] = 1.
    return CMn * (CMp / CMpn)


def btchroma_to_fftmat(btchroma, win=75):
    """Flattened, fft-shifted |fft2| of every 12 x win patch (ftm2d.py:120-138)."""
    nchrm, nbeats = btchroma.shape
    assert nchrm == 12, 'beat-aligned matrix transposed?'
    if nbeats < win:
        return None
    fftmat = np.zeros((nchrm * win, nbeats - win + 1))
    for i in range(nbeats - win + 1):
        # scipy.fftpack keeps float32 input in single precision (complex64), as in the reference
        F = scipy.fftpack.fft2(btchroma[:, i:i + win])
        F = np.sqrt(np.real(F) ** 2 + np.imag(F) ** 2)
        fftmat[:, i] = scipy.fftpack.fftshift(F).flatten()
    return fftmat


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

    def load_features(self, i, do_plot=False):
        filepath = "%s_%i.h5" % (self.get_cacheprefix(), i)
        if i in self.shingles:
            return self.shingles[i]
        elif os.path.exists(filepath):
            self.shingles[i] = featio.load(filepath)['shingle']
            CoverAlgorithm.load_features(self, i)      # clique info as a side effect
            return self.shingles[i]
        feats = CoverAlgorithm.load_features(self, i)
        hpcp_orig = np.asarray(feats[self.chroma_type]).T
        onsets = feats['madmom_features']['onsets']
        hpcp = _sync_median(hpcp_orig, onsets)
        chroma = chrompwr(hpcp, self.PWR)
        shingles = btchroma_to_fftmat(chroma, self.WIN).T
        Norm = np.sqrt(np.sum(shingles ** 2, 1))
        Norm[Norm == 0] = 1
        shingles = np.log(self.C * shingles / Norm[:, None] + 1)
        shingle = np.median(shingles, 0)
        shingle = shingle / np.sqrt(np.sum(shingle ** 2))
        self.shingles[i] = shingle
        featio.save(filepath, {'shingle': shingle})
        return shingle

    def similarity(self, idxs):
        idxs = np.asarray(idxs).reshape(-1, 2)
        if len(idxs) == 0:
            return
        if self._D is None:
            S = np.stack([np.asarray(self.load_features(i), dtype=np.float64) for i in range(self.N)])
            self._D = self.context().ftm2d_allpairs(S)
        self.Ds['main'][idxs[:, 0], idxs[:, 1]] = self._D[idxs[:, 0], idxs[:, 1]]
