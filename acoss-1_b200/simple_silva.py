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


def _hann_window(n):
    """scipy.signal.get_window('hann', n, fftbins=False) == librosa.filters.get_window (symmetric Hann)."""
    k = np.arange(n)
    return 0.5 - 0.5 * np.cos(2.0 * np.pi * k / (n - 1))


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
        if i in self.all_feats:
            return self.all_feats[i]
        feats = CoverAlgorithm.load_features(self, i)
        feat_orig = np.asarray(feats[self.chroma_type]).T
        new_feat = np.zeros((feat_orig.shape[0], int(feat_orig.shape[1] / self.SKIP)))
        for k in range(new_feat.shape[1]):
            new_feat[:, k] = np.mean(feat_orig[:, k * self.SKIP:k * self.SKIP + self.WIN], axis=1)
        self.all_feats[i] = self.smooth(new_feat)
        return self.all_feats[i]

    def smooth(self, feat, win_len_smooth=4):
        """Hann smoothing along time ('same', zero fill) then L2-normalise every column
        (simple_silva.py:56-66; librosa.util.normalize(norm=2, axis=0) leaves tiny columns as is)."""
        win = _hann_window(win_len_smooth + 2)
        win = win / np.sum(win)
        out = np.stack([np.convolve(row, win, mode='full') for row in feat])
        start = (len(win) - 1) // 2
        out = out[:, start:start + feat.shape[1]]
        length = np.sqrt(np.sum(np.abs(out) ** 2, axis=0, keepdims=True))
        length[length < np.finfo(out.dtype).tiny] = 1.0
        return out / length

    def oti(self, seq_a, seq_b):
        """Host restatement of simple_silva.py:45-54 (kept for API parity; `similarity` does it on the device)."""
        profile_a = np.sum(seq_a, 1)
        profile_b = np.sum(seq_b, 1)
        oti_vec = np.zeros(12)
        for i in range(12):
            oti_vec[i] = np.dot(profile_a, np.roll(profile_b, i, axis=0))
        sorted_index = np.argsort(oti_vec)
        return np.roll(seq_b, sorted_index[-1], axis=0), sorted_index

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
