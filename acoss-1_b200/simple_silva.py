# -*- coding: utf-8 -*-
"""
Silva, D. F., Yeh, C. C. M., Zhu, Y., Batista, G., & Keogh, E. (2018). Fast Similarity Matrix
Profile for Music Analysis and Exploration. IEEE Transactions on Multimedia.

Host-side mirror of the reference's `Simple` (acoss/algorithms/simple_silva.py:26-126): same
constructor keywords, `load_features`, `oti`, `smooth`, `similarity(idxs)`.  The per-pair OTI +
matrix-profile join + median (:45-54, :68-118) run on the GPU (acoss_simple_pairs), all ordered
pairs of a `similarity` call in one batch; so does the per-track preprocessing of `load_features` /
`smooth` (:34-43, :56-66) when the tracks are loaded in bulk (`preload`, acoss_simple_preprocess).  Features are cached (the reference re-reads both
files for every pair; results are unaffected).
"""
import numpy as np

from .algorithm_template import CoverAlgorithm

__all__ = ['Simple']


class Simple(CoverAlgorithm):
    """
    Attributes
    ----------
    Same as CoverAlgorithms, plus
    SSLEN=10, the length of subsequence used by SiMPle
    WIN=200, the window length for the dimensionality reduction
    SKIP=100, how many frames the dim reduction will skip each step
    """

    def __init__(self, dataset_csv, datapath, chroma_type='hpcp', shortname='Covers80',
                 SSLEN=10, WIN=200, SKIP=100, cachedir="cache", device=None):
        self.SSLEN = SSLEN
        self.WIN = WIN
        self.SKIP = SKIP
        self.chroma_type = chroma_type
        self.all_feats = {}
        self._uploaded = False
        CoverAlgorithm.__init__(self, dataset_csv=dataset_csv, name="SiMPle", datapath=datapath,
                                shortname=shortname, cachedir=cachedir, device=device)

    def load_features(self, i, do_plot=False):
        """(12, int(T / SKIP)) float64 features of song i (simple_silva.py:34-43): WIN-frame window means every SKIP frames,
        Hann smoothing and column L2 normalisation run on the device (acoss_simple_preprocess); cached."""
        if i not in self.all_feats:
            feats = CoverAlgorithm.load_features(self, i)
            raw = np.asarray(feats[self.chroma_type], np.float32)
            self.all_feats[i] = self.context().simple_preprocess([raw], win=self.WIN, skip=self.SKIP)[0]
        return self.all_feats[i]

    def smooth(self, feat, win_len_smooth=4):
        """Hann smoothing along time ('same', zero fill) then L2-normalise every column (simple_silva.py:56-66), on the device:
        the smoothing stage of acoss_simple_preprocess applied to already-reduced (12, L) features (WIN = SKIP = 1)."""
        feat = np.asarray(feat, np.float64)
        return self.context().simple_preprocess([np.ascontiguousarray(feat.T, np.float32)], win=1, skip=1,
                                                win_len_smooth=win_len_smooth)[0]

    def oti(self, seq_a, seq_b):
        """Simple.oti (simple_silva.py:45-54): roll seq_b's 12 bins by the shift that maximises the dot product of the two
        time-summed profiles (np.argsort, last index).  Public helper of the reference class; `similarity` performs the same
        step inside the pair kernel."""
        prof_a, prof_b = np.sum(seq_a, 1), np.sum(seq_b, 1)
        order = np.argsort([np.dot(prof_a, np.roll(prof_b, k, axis=0)) for k in range(12)])
        return np.roll(seq_b, order[-1], axis=0), order

    def preload(self):
        """All tracks at once: WIN/SKIP means, Hann smoothing and column normalisation run on the device
        (acoss_simple_preprocess) for every track `load_features` has not produced yet, then the features are
        uploaded for the pair kernel."""
        if self._uploaded:
            return
        missing = [i for i in range(self.N) if i not in self.all_feats]
        if missing:
            raw = [np.asarray(CoverAlgorithm.load_features(self, i)[self.chroma_type], np.float32) for i in missing]
            feats = self.context().simple_preprocess(raw, win=self.WIN, skip=self.SKIP)
            for i, f in zip(missing, feats):
                self.all_feats[i] = f
        self.context().simple_upload([self.all_feats[i] for i in range(self.N)])
        self._uploaded = True

    def pair_cost(self, idxs):
        self.preload()
        n = np.array([self.all_feats[i].shape[1] for i in range(self.N)], dtype=np.int64)
        return n[idxs[:, 0]] * n[idxs[:, 1]]

    def similarity(self, idxs):
        idxs = np.asarray(idxs).reshape(-1, 2)
        if len(idxs) == 0:
            return
        self.preload()
        sim = self.context().simple_pairs(idxs, self.SSLEN)
        self.Ds['main'][idxs[:, 0], idxs[:, 1]] = sim
