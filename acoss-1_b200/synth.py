# -*- coding: utf-8 -*-
"""
Synthetic cover-song datasets: there is no network for
covers80 / Da-TACOS, so benchmarks and parity tests run on seeded synthetic HPCP
of the same shape.

Clique model: each work has a latent chroma sequence (piecewise-constant chord
path over 12 bins); each cover = that sequence with a random tempo warp (+-20 %),
a random key shift (roll 0..11), an optional section deletion and additive noise;
every frame is max-normalised to 1 like HPCP.  Singleton "noise" tracks are
independent works.  Track lengths are jittered so that no two tracks share a
length (this breaks score ties after normalize_by_length).

Files are written in the reference's layout `feature_dir/work_id/track_id.h5`
(dict with keys label / hpcp, see featio.py) + a `work_id,track_id` csv.
"""
import os

import numpy as np

from . import featio

# (n_works, tracks_per_work, n_singletons)
SHAPES = {
    "covers80": (80, 2, 0),            # C1: 160 tracks
    "datacos1000": (67, 13, 129),      # C2: 67 x 13 + 129 singletons = 1000
    "datacos15000": (1000, 13, 2000),  # C5
}
# raw length range and the downsample_fac that goes with the regime
REGIMES = {
    "S": dict(len_lo=1600, len_hi=2400, fac=40),     # as written: n ~ 40-60 after /40
    "R": dict(len_lo=15000, len_hi=25000, fac=40),   # real Da-TACOS shape at 86 fps: n ~ 375-625
    "L": dict(len_lo=1600, len_hi=2400, fac=1),      # DP-meaningful: n ~ 1600-2400 at the DP stage
}


def _latent(rng, length):
    """piecewise-constant chord path, (length, 12) float32 in [0, 1]"""
    out = np.empty((length, 12), np.float32)
    t = 0
    while t < length:
        dur = int(rng.integers(8, 41))
        root = int(rng.integers(0, 12))
        third = 4 if rng.random() < 0.6 else 3
        tpl = np.full(12, 0.05, np.float32)
        tpl[root] = 1.0
        tpl[(root + third) % 12] = 0.8
        tpl[(root + 7) % 12] = 0.9
        tpl[(root + 10) % 12] += 0.2 * rng.random()
        out[t:t + dur] = tpl
        t += dur
    return out


# the "hard" cover model (bench.py C1/C2, tests/golden/serra09_L48.npz): heavier noise and a fraction of every cover's chord
# segments replaced by unrelated chords, so that retrieval is NOT saturated (MAP < 1, mean rank > 1) and an evaluation
# tuple that differs from the oracle's can actually be seen
HARD = dict(noise=0.25, p_sub=0.25)


def _render(rng, latent, length, noise, p_sub=0.0):
    """warp latent to `length` frames, key shift, optional deletion, chord substitutions, noise, max-normalise"""
    L0 = latent.shape[0]
    # piecewise tempo: 8 segments with speeds in [0.8, 1.2]
    speeds = rng.uniform(0.8, 1.2, size=8)
    pos = np.cumsum(np.repeat(speeds, int(np.ceil(length / 8.0)))[:length])
    pos = pos / pos[-1] * (L0 - 1)
    x = latent[np.floor(pos).astype(np.int64)]
    if rng.random() < 0.5 and length > 200:       # drop a section, re-stretch by repetition of the tail
        a = int(rng.integers(0, length // 2))
        b = a + int(rng.integers(length // 20, length // 7))
        x = np.concatenate([x[:a], x[b:], x[a:b][::-1][: b - a]], axis=0)[:length]
    if p_sub > 0.0:                               # replace a fraction of the chord segments by unrelated chords
        x = np.array(x)
        t = 0
        while t < length:
            dur = int(rng.integers(8, 41))
            if rng.random() < p_sub:
                x[t:t + dur] = _latent(rng, dur)[:len(x[t:t + dur])]
            t += dur
    x = np.roll(x, int(rng.integers(0, 12)), axis=1)
    x = x + np.abs(rng.normal(0.0, noise, size=x.shape)).astype(np.float32)
    x = x / np.maximum(x.max(axis=1, keepdims=True), 1e-9)
    return np.ascontiguousarray(x, dtype=np.float32)


def make_tracks(n_works, per_work, n_singletons, seed, len_lo, len_hi, noise=0.12, p_sub=0.0):
    """Returns (labels: list[str], work_ids: list[str], track_ids: list[str], hpcps: list[(T,12) f32])."""
    rng = np.random.default_rng(seed)
    n = n_works * per_work + n_singletons
    span = len_hi - len_lo + 1
    if span >= n:
        lengths = len_lo + rng.choice(span, size=n, replace=False)
    else:                                        # more tracks than distinct lengths in range
        lengths = len_lo + rng.integers(0, span, size=n)
    labels, works, tids, hp = [], [], [], []
    k = 0
    for w in range(n_works + n_singletons):
        cnt = per_work if w < n_works else 1
        lat = _latent(rng, int(rng.integers(len_lo, len_hi + 1)))
        for c in range(cnt):
            labels.append("W_%d" % w)
            works.append("W_%d" % w)
            tids.append("P_%d_%d" % (w, c))
            hp.append(_render(rng, lat, int(lengths[k]), noise, p_sub))
            k += 1
    return labels, works, tids, hp


def make_named(shape="covers80", regime="S", seed=None, scale=1.0, hard=False, **model):
    """Tracks of one of SHAPES x REGIMES.  `scale` < 1 shrinks the number of works (tests); `hard` selects the HARD
    cover model; other keywords (noise, p_sub) override the cover model."""
    nw, pw, ns = SHAPES[shape]
    if scale != 1.0:
        nw, ns = max(1, int(round(nw * scale))), int(round(ns * scale))
    rg = REGIMES[regime]
    if seed is None:
        seed = {"covers80": 80, "datacos1000": 1000, "datacos15000": 15000}[shape]
    kw = dict(HARD) if hard else {}
    kw.update(model)
    labels, works, tids, hp = make_tracks(nw, pw, ns, seed, rg["len_lo"], rg["len_hi"], **kw)
    return dict(labels=labels, work_ids=works, track_ids=tids, hpcp=hp, fac=rg["fac"])


def write_dataset(feature_dir, ds, csv_path, extra=None):
    """feature_dir/work_id/track_id.h5 + csv.  feature_dir gets its trailing '/' (SURVEY Q14)."""
    if not feature_dir.endswith("/"):
        feature_dir += "/"
    os.makedirs(feature_dir, exist_ok=True)
    with open(csv_path, "w") as f:
        f.write("work_id,track_id\n")
        for i, (w, t) in enumerate(zip(ds["work_ids"], ds["track_ids"])):
            os.makedirs(os.path.join(feature_dir, w), exist_ok=True)
            d = {"label": ds["labels"][i], "hpcp": ds["hpcp"][i]}
            if extra:
                d.update({k: v[i] for k, v in extra.items()})
            featio.save(os.path.join(feature_dir, w, t + ".h5"), d)
            f.write("%s,%s\n" % (w, t))
    return feature_dir


def make_block_features(n_works, per_work, n_singletons, seed, nb_lo, nb_hi, n_mfcc=20, noise=0.35):
    """Synthetic EarlyFusion block features (what EarlyFusion.load_features returns,
    acoss/algorithms/earlyfusion_traile.py:67-154): per track a dict with
    'mfccs' (n_blocks, 50*n_mfcc), 'ssms' (n_blocks, 1225), 'chromas' (n_blocks, 480) float32 and
    'chroma_med' (12,) float32.  Covers of a work share a latent block sequence (random start offset,
    additive noise, random key shift of the chroma blocks)."""
    rng = np.random.default_rng(seed)
    d_m, d_s, d_c = 50 * n_mfcc, 1225, 480
    labels, works, tids, feats = [], [], [], []

    def render(lat, nb, start, key):
        m = lat["m"][start:start + nb] + noise * rng.standard_normal((nb, d_m))
        m = m - m.mean(1, keepdims=True)
        m = m / np.sqrt((m ** 2).sum(1, keepdims=True))
        s = np.abs(lat["s"][start:start + nb] + noise * rng.standard_normal((nb, d_s)))
        c = np.abs(lat["c"][start:start + nb] + 0.5 * noise * rng.standard_normal((nb, d_c)))
        c = np.roll(c.reshape(nb, d_c // 12, 12), key, axis=2).reshape(nb, d_c)
        med = np.roll(lat["med"] + 0.02 * rng.random(12), key)
        return {"mfccs": m.astype(np.float32), "ssms": s.astype(np.float32), "chromas": c.astype(np.float32),
                "chroma_med": med.astype(np.float32)}

    def latent(length):
        # slowly varying paths: neighbouring blocks overlap by 19 of 20 beats in the real features
        def path(d, scale):
            x = np.cumsum(rng.standard_normal((length, d)), 0) * scale
            return x - x.mean(0, keepdims=True)
        return {"m": path(d_m, 0.25), "s": path(d_s, 0.25) + 1.0, "c": np.abs(path(d_c, 0.2)) + 0.1,
                "med": rng.random(12) + 0.2}

    t = 0
    for w in range(n_works + n_singletons):
        n_cov = per_work if w < n_works else 1
        lat = latent(nb_hi + 16)
        for c in range(n_cov):
            nb = int(rng.integers(nb_lo, nb_hi + 1))
            start = int(rng.integers(0, 16))
            key = int(rng.integers(0, 12))
            feats.append(render(lat, nb, start, key))
            labels.append("W_%d" % w)
            works.append("W_%d" % w)
            tids.append("P_%d" % t)
            t += 1
    return dict(labels=labels, work_ids=works, track_ids=tids, block_feats=feats)
