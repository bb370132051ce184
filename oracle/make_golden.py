# -*- coding: utf-8 -*-
"""
CPU ORACLE (golden generator) -- TEST INFRASTRUCTURE ONLY, THIS CONTAINER ONLY.

Freezes outputs of the reference's own Python code (imported verbatim from
/root/reference through oracle/refload.py) and of the oracle under tests/golden/,
so that the GPU box -- where /root/reference does not exist -- can check against
them.  Run:  python -m oracle.make_golden        (from the repository root)

Files written (all small .npz):
  sw_reference.npz       smith_waterman_constrained (numba, verbatim) on random binary matrices
  eval_reference.npz     CoverAlgorithm.getEvalStatistics (verbatim) on random + tied matrices
  snf_reference.npz      doSimilarityFusion (verbatim) at (n_mats, N, K, niters) of SURVEY App. B-8
  serra09_small.npz      16-track synthetic set: reference Serra09 / ChenFusion classes run
                         verbatim on top of the C restatement of essentia (parity UNPINNED for
                         that part) -> Ds before/after normalisation, eval tuples, a few CRPs
  percentile_kat.npz     oracle percentile vs np.percentile(method='linear') in float32
  earlyfusion_large.npz  EarlyFusion pairs at BASELINE C3 size (300-600 blocks, n_mfcc = 20): the reference's free
                         functions + class run verbatim; bit-packed binary matrices, sub-sampled CSM rows, the
                         k-th / (k+1)-th neighbour gaps of every row (near-tie diagnostics), all scores
  simple_large.npz       SiMPle at BASELINE C5 size (150-250 columns): 56 ordered pairs, float64 scores
  serra09_L48.npz        48 tracks at the L length regime (1128 pairs, M, N ~ 1600-2400): oracle scores
                         (threaded C restatement), then the reference class's normalize_by_length +
                         getEvalStatistics verbatim on that matrix; harder synthetic covers (MAP < 1)
"""
import contextlib
import io
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
GOLD = os.path.join(ROOT, "tests", "golden")


def quiet(fn, *a, **k):
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        return fn(*a, **k)


def gen_sw(ns):
    rng = np.random.default_rng(1234)
    shapes = [(3, 9), (4, 4), (5, 7), (12, 12), (40, 37), (64, 64), (100, 130), (257, 190), (300, 512)]
    mats, scores = [], []
    for (M, N) in shapes:
        for dens in (0.05, 0.15, 0.5):
            B = (rng.random((M, N)) < dens).astype(np.uint8)
            # add a few diagonal runs so that scores are non-trivial
            for _ in range(3):
                L = int(rng.integers(2, max(3, min(M, N))))
                i0, j0 = int(rng.integers(0, M - 1)), int(rng.integers(0, N - 1))
                for k in range(L):
                    if i0 + k < M and j0 + k < N:
                        B[i0 + k, j0 + k] = 1
            mats.append(B)
            scores.append(float(ns.align.smith_waterman_constrained(B)))
    d = {"n": len(mats), "scores": np.array(scores, np.float64)}
    for k, B in enumerate(mats):
        d["B%d" % k] = B
    np.savez_compressed(os.path.join(GOLD, "sw_reference.npz"), **d)
    print("sw_reference:", len(mats), "matrices")


def _eval_verbatim(ns, D, labels, tmp):
    """Run the reference's getEvalStatistics on matrix D with clique labels."""
    csv = os.path.join(tmp, "e.csv")
    with open(csv, "w") as f:
        f.write("work_id,track_id\n")
        for i, l in enumerate(labels):
            f.write("%s,t%d\n" % (l, i))
    cwd = os.getcwd()
    os.chdir(tmp)
    try:
        obj = quiet(ns.template.CoverAlgorithm, csv, name="E", datapath=tmp + "/", shortname="g", cachedir=os.path.join(tmp, "cache"))
        for i, l in enumerate(labels):
            obj.cliques.setdefault(l, set([])).add(i)
        obj.Ds["main"][:, :] = D
        res = quiet(obj.getEvalStatistics, "main")
    finally:
        os.chdir(cwd)
    return res


def gen_eval(ns):
    rng = np.random.default_rng(99)
    out = {}
    cases = []
    # (n_works, per_work, singletons)
    for ci, (nw, pw, nsng, quant) in enumerate([(8, 2, 0, False), (5, 4, 7, False), (6, 13, 12, False), (10, 3, 5, True)]):
        labels = []
        for w in range(nw):
            labels += ["W%d" % w] * pw
        labels += ["S%d" % s for s in range(nsng)]
        perm = rng.permutation(len(labels))
        labels = [labels[p] for p in perm]
        n = len(labels)
        same = np.array([[labels[i] == labels[j] for j in range(n)] for i in range(n)])
        D = rng.random((n, n)).astype(np.float32) + 0.35 * same
        if quant:   # heavy quantisation -> ties exist; only tie-free-robust fields are compared
            D = np.round(D * 4) / 4
        D = D.astype(np.float32)
        with tempfile.TemporaryDirectory() as tmp:
            MR, MRR, MDR, MAP, tops = _eval_verbatim(ns, D, labels, tmp)
        out["D%d" % ci] = D
        out["labels%d" % ci] = np.array(labels)
        out["res%d" % ci] = np.array([MR, MRR, MDR, MAP] + list(tops), np.float64)
        out["ties%d" % ci] = np.array(int(quant))
        cases.append(ci)
    out["n"] = np.array(len(cases))
    np.savez_compressed(os.path.join(GOLD, "eval_reference.npz"), **out)
    print("eval_reference:", len(cases), "cases")


def gen_snf(ns):
    rng = np.random.default_rng(7)
    out = {}
    cfgs = [(2, 64, 5, 3), (3, 200, 20, 20), (4, 120, 20, 20)]
    for ci, (nm, N, K, nit) in enumerate(cfgs):
        mats = []
        for _ in range(nm):
            X = rng.random((N, 6))
            Dm = np.sqrt(((X[:, None, :] - X[None, :, :]) ** 2).sum(-1)) + 0.05 * rng.random((N, N))
            mats.append(Dm.astype(np.float32))
        Ws, fused = ns.snf.doSimilarityFusion([np.array(m) for m in mats], K=K, niters=nit, reg_diag=1)
        out["cfg%d" % ci] = np.array([nm, N, K, nit])
        for k, m in enumerate(mats):
            out["D%d_%d" % (ci, k)] = m
        out["fused%d" % ci] = fused
    out["n"] = np.array(len(cfgs))
    np.savez_compressed(os.path.join(GOLD, "snf_reference.npz"), **out)
    print("snf_reference:", len(cfgs), "configs")


def gen_serra(ns):
    import acoss_b200
    from acoss_b200 import synth
    from oracle import oracle as O
    ds = synth.make_named("covers80", "S", seed=4242, scale=0.15)    # 12 works x 2 = 24 tracks, T ~ 1600-2400
    out = {"n_tracks": np.array(len(ds["hpcp"])), "labels": np.array(ds["labels"]), "fac": np.array(ds["fac"])}
    for i, h in enumerate(ds["hpcp"]):
        out["hpcp%d" % i] = h
    with tempfile.TemporaryDirectory() as tmp:
        fdir = synth.write_dataset(os.path.join(tmp, "feat"), ds, os.path.join(tmp, "ds.csv"))
        cwd = os.getcwd()
        os.chdir(tmp)
        try:
            # --- reference Serra09 verbatim (orchestration) on the essentia restatement
            s = quiet(ns.serra09.Serra09, os.path.join(tmp, "ds.csv"), fdir, shortname="g")
            quiet(s.all_pairwise, 0, 1, True)
            out["serra_raw"] = np.array(s.Ds["main"])
            quiet(s.normalize_by_length)
            out["serra_norm"] = np.array(s.Ds["main"])
            out["serra_eval"] = np.array([x for x in quiet(s.getEvalStatistics, "main")[:4]], np.float64)
            out["lengths"] = np.array([s.all_feats[i].shape[0] for i in range(s.N)], np.int32)
            for i in range(s.N):
                out["ds%d" % i] = np.asarray(s.all_feats[i], np.float32)
            # --- reference ChenFusion verbatim
            c = quiet(ns.chen.ChenFusion, os.path.join(tmp, "ds.csv"), fdir, shortname="g")
            quiet(c.all_pairwise, 0, 1, True)
            out["chen_q_raw"] = np.array(c.Ds["qmax"])
            out["chen_d_raw"] = np.array(c.Ds["dmax"])
            with np.errstate(divide="ignore"):
                quiet(c.normalize_by_length)
            out["chen_q_norm"] = np.array(c.Ds["qmax"])
            out["chen_d_norm"] = np.array(c.Ds["dmax"])
            with np.errstate(all="ignore"):
                quiet(c.do_late_fusion)
            out["chen_late"] = np.array(c.Ds["Late"])
            ev = {}
            for k in ("qmax", "dmax", "Late"):
                ev[k] = quiet(c.getEvalStatistics, k)
                out["chen_eval_" + k] = np.array(list(ev[k][:4]), np.float64)
        finally:
            os.chdir(cwd)
    # a few full CRPs + thresholds from the oracle
    pairs = [(0, 1), (2, 7), (5, 4), (10, 3), (23, 22)]
    out["crp_pairs"] = np.array(pairs, np.int32)
    for k, (a, b) in enumerate(pairs):
        R, info = O.chroma_cross_similarity(out["ds%d" % a], out["ds%d" % b], full=True)
        out["crp%d" % k] = R
        out["thrq%d" % k] = info["thr_q"]
        out["thrr%d" % k] = info["thr_r"]
        out["oti%d" % k] = np.array(info["oti"])
    np.savez_compressed(os.path.join(GOLD, "serra09_small.npz"), **out)
    print("serra09_small: MAP", out["serra_eval"][3], "chen late MAP", out["chen_eval_Late"][3])



def simple_ftm_dataset():
    """Deterministic small dataset for the SiMPle / FTM2D goldens: 5 works x 2 covers, raw T ~ 2400-3600
    (SiMPle: 24-36 columns; FTM2D: beats every ~18 frames -> 130-200 beats >= WIN = 75)."""
    from acoss_b200 import synth
    labels, works, tids, hp = synth.make_tracks(5, 2, 0, 777, 2400, 3600)
    rng = np.random.default_rng(778)
    onsets = []
    for h in hp:
        gaps = rng.integers(14, 23, size=h.shape[0] // 14 + 2)
        o = np.cumsum(gaps)
        onsets.append(o[o < h.shape[0]].astype(np.int64))
    ds = dict(labels=labels, work_ids=works, track_ids=tids, hpcp=hp, fac=1)
    extra = {"madmom_features": [{"onsets": o} for o in onsets]}
    return ds, extra


def gen_simple_ftm(ns):
    from acoss_b200 import synth
    ds, extra = simple_ftm_dataset()
    out = {"hpcp_sum": np.array([float(np.sum(h, dtype=np.float64)) for h in ds["hpcp"]])}
    with tempfile.TemporaryDirectory() as tmp:
        fdir = synth.write_dataset(os.path.join(tmp, "feat"), ds, os.path.join(tmp, "ds.csv"), extra=extra)
        cwd = os.getcwd()
        os.chdir(tmp)
        try:
            s = quiet(ns.simple.Simple, os.path.join(tmp, "ds.csv"), fdir, shortname="g")
            for i in range(s.N):
                out["simple_feat%d" % i] = np.asarray(quiet(s.load_features, i), np.float64)
            quiet(s.all_pairwise, 0, 1, False)
            out["simple_D"] = np.array(s.Ds["main"])
            # float64 scores of a few ordered pairs straight from simple_sim (Ds is float32)
            prs = [(0, 1), (1, 0), (2, 7), (9, 4), (3, 3)]
            out["simple_pairs"] = np.array(prs, np.int32)
            out["simple_scores"] = np.array([-s.simple_sim(out["simple_feat%d" % a], s.oti(out["simple_feat%d" % a], out["simple_feat%d" % b])[0])
                                             for a, b in prs], np.float64)
            out["simple_eval"] = np.array(list(quiet(s.getEvalStatistics, "main")[:4]), np.float64)
            f = quiet(ns.ftm2d.FTM2D, os.path.join(tmp, "ds.csv"), fdir, shortname="g")
            for i in range(f.N):
                out["ftm_shingle%d" % i] = np.asarray(quiet(f.load_features, i), np.float64)
            quiet(f.all_pairwise, 0, 1, True)
            out["ftm_D"] = np.array(f.Ds["main"])
            out["ftm_eval"] = np.array(list(quiet(f.getEvalStatistics, "main")[:4]), np.float64)
        finally:
            os.chdir(cwd)
    np.savez_compressed(os.path.join(GOLD, "simple_ftm2d_reference.npz"), **out)
    print("simple_ftm2d_reference: simple MAP", out["simple_eval"][3], "ftm MAP", out["ftm_eval"][3])


def early_dataset():
    """Deterministic block features for the EarlyFusion goldens: 11 works x 2 + 2 singletons = 24 tracks
    (K + 1 = 21 < N_t for the N x N fusion), 36-60 blocks each."""
    from acoss_b200 import synth
    ds = synth.make_block_features(11, 2, 2, 991, 36, 60)
    ds["hpcp"] = [np.zeros((4, 12), np.float32) for _ in ds["labels"]]
    ds["fac"] = 1
    return ds


def gen_early(ns):
    from acoss_b200 import synth
    ds = early_dataset()
    out = {"checksum": np.array([float(np.sum(f["mfccs"], dtype=np.float64) + np.sum(f["chromas"], dtype=np.float64))
                                 for f in ds["block_feats"]])}
    with tempfile.TemporaryDirectory() as tmp:
        fdir = synth.write_dataset(os.path.join(tmp, "feat"), ds, os.path.join(tmp, "ds.csv"))
        cwd = os.getcwd()
        os.chdir(tmp)
        try:
            e = quiet(ns.early.EarlyFusion, os.path.join(tmp, "ds.csv"), fdir, shortname="g")
            for i, f in enumerate(ds["block_feats"]):
                e.all_block_feats[i] = f
            quiet(e.all_pairwise, 0, 1, True)
            quiet(e.get_all_clique_ids)       # the injected block features bypass load_features' side effect
            for s_ in ("mfccs", "ssms", "chromas", "early"):
                out["D_" + s_] = np.array(e.Ds[s_])
            # matrices of two pairs, straight from the reference's free functions
            prs = [(0, 1), (5, 12)]
            out["mat_pairs"] = np.array(prs, np.int32)
            for k, (a, b) in enumerate(prs):
                f1, f2 = ds["block_feats"][a], ds["block_feats"][b]
                C = [ns.xrec.get_csm(f1["mfccs"], f2["mfccs"]), ns.xrec.get_csm(f1["ssms"], f2["ssms"]),
                     ns.xrec.get_csm_blocked_oti(f1["chromas"], f2["chromas"], f1["chroma_med"], f2["chroma_med"], ns.xrec.get_csm_cosine)]
                Ws = [ns.snf.getWCSM(c, e.K, e.K) for c in C]
                Wsum = np.zeros_like(C[0])
                for w in Ws:
                    Wsum += w
                C.append(np.exp(-Wsum))
                out["csm%d" % k] = np.stack(C).astype(np.float32)
                out["bin%d" % k] = np.stack([np.asarray(ns.xrec.csm_to_binary(c, e.kappa).toarray() if hasattr(ns.xrec.csm_to_binary(c, e.kappa), "toarray")
                                                        else ns.xrec.csm_to_binary(c, e.kappa)) for c in C]).astype(np.uint8)
                out["oti%d" % k] = np.array(ns.xrec.get_oti(f1["chroma_med"], f2["chroma_med"]))
            with np.errstate(all="ignore"):
                quiet(e.do_late_fusion)
            out["D_late"] = np.array(e.Ds["late"])
            out["D_early+late"] = np.array(e.Ds["early+late"])
            for s_ in ("mfccs", "ssms", "chromas", "early", "late", "early+late"):
                out["eval_" + s_] = np.array(list(quiet(e.getEvalStatistics, s_)[:4]), np.float64)
        finally:
            os.chdir(cwd)
    np.savez_compressed(os.path.join(GOLD, "earlyfusion_reference.npz"), **out)
    print("earlyfusion_reference: MAP", {s_: float(out["eval_" + s_][3]) for s_ in ("mfccs", "ssms", "chromas", "early", "late", "early+late")})

def early_large_dataset():
    """Block features at BASELINE C3 size: 2 works x 2 covers + 1 singleton, 300-600 blocks, d = 1000 / 1225 / 480.
    Regenerated from the seed on the GPU box (5 MB per track is too much to commit); `checksum` guards it."""
    from acoss_b200 import synth
    ds = synth.make_block_features(2, 2, 1, 3301, 300, 600, noise=2.0)
    ds["hpcp"] = [np.zeros((4, 12), np.float32) for _ in ds["labels"]]
    ds["fac"] = 1
    return ds


EARLY_LARGE_PAIRS = [(0, 1), (2, 3), (1, 4), (0, 3)]


def gen_early_large(ns):
    from acoss_b200 import synth
    ds = early_large_dataset()
    bf = ds["block_feats"]
    out = {"checksum": np.array([float(np.sum(f["mfccs"], dtype=np.float64) + np.sum(f["ssms"], dtype=np.float64) +
                                       np.sum(f["chromas"], dtype=np.float64)) for f in bf]),
           "n_blocks": np.array([f["mfccs"].shape[0] for f in bf], np.int32),
           "mat_pairs": np.array(EARLY_LARGE_PAIRS, np.int32)}
    with tempfile.TemporaryDirectory() as tmp:
        fdir = synth.write_dataset(os.path.join(tmp, "feat"), ds, os.path.join(tmp, "ds.csv"))
        cwd = os.getcwd()
        os.chdir(tmp)
        try:
            e = quiet(ns.early.EarlyFusion, os.path.join(tmp, "ds.csv"), fdir, shortname="gl")
            for i, f in enumerate(bf):
                e.all_block_feats[i] = f
            quiet(e.all_pairwise, 0, 1, True)
            for s_ in ("mfccs", "ssms", "chromas", "early"):
                out["D_" + s_] = np.array(e.Ds[s_])
            for k, (a, b) in enumerate(EARLY_LARGE_PAIRS):
                f1, f2 = bf[a], bf[b]
                C = [ns.xrec.get_csm(f1["mfccs"], f2["mfccs"]), ns.xrec.get_csm(f1["ssms"], f2["ssms"]),
                     ns.xrec.get_csm_blocked_oti(f1["chromas"], f2["chromas"], f1["chroma_med"], f2["chroma_med"], ns.xrec.get_csm_cosine)]
                Ws = [ns.snf.getWCSM(c, e.K, e.K) for c in C]
                Wsum = np.zeros_like(C[0])
                for w in Ws:
                    Wsum += w
                C.append(np.exp(-Wsum))
                C = [np.asarray(c, np.float32) for c in C]
                M, N = C[0].shape
                kk = int(np.round(e.kappa * N))
                bins, gaps = [], []
                for c in C:
                    B = ns.xrec.csm_to_binary(c, e.kappa)
                    B = np.asarray(B.toarray() if hasattr(B, "toarray") else B).astype(np.uint8)
                    bins.append(np.packbits(B, axis=1))
                    srt = np.sort(c.astype(np.float64), axis=1)
                    gaps.append(np.stack([srt[:, kk - 1], srt[:, kk]], 1))       # k-th and (k+1)-th smallest of every row
                out["csm_rows%d" % k] = np.stack([c[::16] for c in C])
                out["binpk%d" % k] = np.stack(bins)
                out["kth%d" % k] = np.stack(gaps)
                out["shape%d" % k] = np.array([M, N, kk], np.int32)
                out["oti%d" % k] = np.array(ns.xrec.get_oti(f1["chroma_med"], f2["chroma_med"]))
                out["scores%d" % k] = np.array([float(ns.align.smith_waterman_constrained(np.unpackbits(bp, axis=1)[:, :N])) for bp in bins])
        finally:
            os.chdir(cwd)
    np.savez_compressed(os.path.join(GOLD, "earlyfusion_large.npz"), **out)
    print("earlyfusion_large:", [tuple(out["shape%d" % k]) for k in range(len(EARLY_LARGE_PAIRS))],
          [out["scores%d" % k].tolist() for k in range(len(EARLY_LARGE_PAIRS))])


def gen_simple_large(ns):
    """SiMPle at BASELINE C5 size: 8 tracks of 150-250 columns (raw T ~ 15-25 k at SKIP = 100), all 56 ordered pairs."""
    from acoss_b200 import synth
    labels, works, tids, hp = synth.make_tracks(3, 2, 2, 5150, 15000, 25000)
    ds = dict(labels=labels, work_ids=works, track_ids=tids, hpcp=hp, fac=1)
    out = {"hpcp_sum": np.array([float(np.sum(h, dtype=np.float64)) for h in hp])}
    with tempfile.TemporaryDirectory() as tmp:
        fdir = synth.write_dataset(os.path.join(tmp, "feat"), ds, os.path.join(tmp, "ds.csv"))
        cwd = os.getcwd()
        os.chdir(tmp)
        try:
            s = quiet(ns.simple.Simple, os.path.join(tmp, "ds.csv"), fdir, shortname="gl")
            feats = [np.asarray(quiet(s.load_features, i), np.float64) for i in range(s.N)]
            prs = [(a, b) for a in range(s.N) for b in range(s.N) if a != b]
            out["pairs"] = np.array(prs, np.int32)
            out["scores"] = np.array([-s.simple_sim(feats[a], s.oti(feats[a], feats[b])[0]) for a, b in prs], np.float64)
            out["n"] = np.array(s.N)
            for i, f in enumerate(feats):
                out["feat%d" % i] = f
        finally:
            os.chdir(cwd)
    np.savez_compressed(os.path.join(GOLD, "simple_large.npz"), **out)
    print("simple_large:", [f.shape[1] for f in feats], "scores", out["scores"][:4])


def serra_L48_dataset():
    """48 tracks at the L regime (n ~ 1600-2400 frames at the DP stage, downsample_fac = 1) with the HARD cover model
    (noise + chord substitutions, the same generator bench.py uses for C2): 12 works x 3 + 12 singletons."""
    from acoss_b200 import synth
    labels, works, tids, hp = synth.make_tracks(12, 3, 12, 4801, 1600, 2400, **synth.HARD)
    return dict(labels=labels, work_ids=works, track_ids=tids, hpcp=hp, fac=1)


def gen_serra_L48(ns):
    from acoss_b200 import synth
    from oracle import oracle as O
    ds = serra_L48_dataset()
    n = len(ds["hpcp"])
    i, j = np.triu_indices(n, 1)
    pairs = np.stack([i, j], 1).astype(np.int32)
    q, d, cells = O.serra09_pairs(ds["hpcp"], pairs, want_dmax=True, n_threads=os.cpu_count())
    out = {"checksum": np.array([float(np.sum(h, dtype=np.float64)) for h in ds["hpcp"]]), "pairs": pairs,
           "qmax": q, "dmax": d, "labels": np.array(ds["labels"])}
    small = dict(ds)
    small["hpcp"] = [h[:12] for h in ds["hpcp"]]      # the class only needs labels + lengths below
    with tempfile.TemporaryDirectory() as tmp:
        fdir = synth.write_dataset(os.path.join(tmp, "feat"), small, os.path.join(tmp, "ds.csv"))
        cwd = os.getcwd()
        os.chdir(tmp)
        try:
            s = quiet(ns.serra09.Serra09, os.path.join(tmp, "ds.csv"), fdir, shortname="gL", downsample_fac=1)
            for k in range(n):
                quiet(s.load_features, k)
                s.all_feats[k] = ds["hpcp"][k]
            D = np.zeros((n, n), np.float32)
            D[pairs[:, 0], pairs[:, 1]] = q
            D += D.T
            s.Ds["main"][:, :] = D
            out["serra_raw"] = D
            quiet(s.normalize_by_length)
            out["serra_norm"] = np.array(s.Ds["main"])
            out["serra_eval"] = np.array([x for x in quiet(s.getEvalStatistics, "main")[:4]], np.float64)
        finally:
            os.chdir(cwd)
    np.savez_compressed(os.path.join(GOLD, "serra09_L48.npz"), **out)
    print("serra09_L48: MR/MRR/MDR/MAP", out["serra_eval"])


def gen_percentile():
    from oracle import oracle as O
    rng = np.random.default_rng(5)
    vs, ps = [], []
    for n in (1, 2, 3, 11, 12, 41, 200, 1999, 2000):
        v = rng.random(n).astype(np.float32) * 3
        vs.append(v)
        ps.append(O.percentile(v, 0.095))
    d = {"n": np.array(len(vs)), "p": np.array(ps, np.float32)}
    for k, v in enumerate(vs):
        d["v%d" % k] = v
    np.savez_compressed(os.path.join(GOLD, "percentile_kat.npz"), **d)
    print("percentile_kat ok")


def main():
    from oracle import refload
    os.makedirs(GOLD, exist_ok=True)
    ns = refload.load()
    only = sys.argv[1:]
    if only:                      # python -m oracle.make_golden early_large simple_large ...
        for name in only:
            fn = globals()["gen_" + name]
            fn() if name == "percentile" else fn(ns)
        return
    gen_percentile()
    gen_sw(ns)
    gen_eval(ns)
    gen_snf(ns)
    gen_serra(ns)
    gen_simple_ftm(ns)
    gen_early(ns)
    gen_early_large(ns)
    gen_simple_large(ns)
    gen_serra_L48(ns)


if __name__ == "__main__":
    main()
