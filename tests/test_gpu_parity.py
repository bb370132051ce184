"""GPU parity tests (-m gpu): the CUDA path, called through the C-ABI, against the oracle on the
same seeded inputs, against the committed golden fixtures, and -- at large sizes -- through
size-independent properties.  Bar: bit-exact for the CRP / thresholds / integer scores and ranks;
normalised float32 matrices bit-equal (same float64 arithmetic); eval tuples identical."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _tracks(rng, lens, noise=0.15):
    out = []
    base = rng.random((max(lens) + 50, 12)).astype(np.float32)
    for k, n in enumerate(lens):
        x = np.roll(base[k % 7: k % 7 + n], k % 12, axis=1) + noise * rng.random((n, 12)).astype(np.float32)
        x = x / x.max(axis=1, keepdims=True)
        out.append(np.ascontiguousarray(x, np.float32))
    return out


# ---- first kernel: Qmax / Dmax on given binary matrices -------------------------------------
def test_dp_known_answers(ctx):
    L = 12
    I = np.eye(L, dtype=np.uint8)
    G = I.copy(); G[6, 6] = 0
    Z = np.zeros((L, L), np.uint8)
    O1 = np.ones((L, L), np.uint8)
    q = ctx.cover_similarity_from_crp([I, G, Z, O1], "serra09")
    d = ctx.cover_similarity_from_crp([I, G, Z, O1], "chen17")
    assert q.tolist() == [10.0, 8.5, 0.0, 10.0]
    assert d.tolist()[1:] == [8.5, 0.0, 14.0]


@pytest.mark.parametrize("shape", [(3, 3), (2, 50), (50, 2), (41, 37), (64, 128), (130, 129), (300, 511), (257, 513),
                                   (600, 1279), (500, 1281), (400, 1536), (300, 2391), (200, 2560), (90, 2700)])
def test_dp_random_vs_oracle(ctx, oracle, shape):
    rng = np.random.default_rng(shape[0] * 7919 + shape[1])
    mats = []
    for dens in (0.03, 0.1, 0.4):
        B = (rng.random(shape) < dens).astype(np.uint8)
        for _ in range(4):   # diagonal runs
            i0, j0 = int(rng.integers(0, shape[0])), int(rng.integers(0, shape[1]))
            L = int(rng.integers(1, 1 + min(shape)))
            for k in range(L):
                if i0 + k < shape[0] and j0 + k < shape[1]:
                    B[i0 + k, j0 + k] = 1
        mats.append(B)
    for a in ("serra09", "chen17"):
        got = ctx.cover_similarity_from_crp(mats, a)
        exp = [oracle.cover_song_similarity(B, a)[1] for B in mats]
        assert got.tolist() == exp, (a, shape)


def test_dp_ragged_batch_and_general_gammas(ctx, oracle):
    rng = np.random.default_rng(5)
    mats = [(rng.random((int(rng.integers(1, 200)), int(rng.integers(1, 400)))) < 0.2).astype(np.uint8) for _ in range(40)]
    for a in ("serra09", "chen17"):
        assert ctx.cover_similarity_from_crp(mats, a).tolist() == [oracle.cover_song_similarity(B, a)[1] for B in mats]
        for (go, ge) in [(0.5, 0.7), (0.25, 1.0)]:   # generic fp32 kernel
            got = ctx.cover_similarity_from_crp(mats, a, go, ge)
            exp = np.array([oracle.cover_song_similarity(B, a, go, ge)[1] for B in mats], np.float32)
            assert np.array_equal(got, exp)


def test_dp_rejects_non_binary(ctx):
    import acoss_b200
    with pytest.raises(acoss_b200.AcossError):
        ctx.cover_similarity_from_crp([np.full((5, 5), 2, np.uint8)])


def test_dp_transpose_invariance_large(ctx):
    """size-independent property at full size: the recurrences are symmetric in (i, j)."""
    rng = np.random.default_rng(8)
    B = (rng.random((2300, 2391)) < 0.08).astype(np.uint8)
    for a in ("serra09", "chen17"):
        s = ctx.cover_similarity_from_crp([B, np.ascontiguousarray(B.T[:, :2300])], a)
        assert s[0] == s[1] and s[0] > 0


# ---- constrained Smith-Waterman ----------------------------------------------------------------
def test_sw_vs_reference_golden(ctx, golden):
    g = golden("sw_reference.npz")
    mats = [g["B%d" % k] for k in range(int(g["n"]))]
    got = ctx.sw_constrained(mats)
    assert np.array_equal(got.astype(np.float32), g["scores"].astype(np.float32))


def test_sw_vs_oracle_random(ctx, oracle):
    rng = np.random.default_rng(21)
    mats = [(rng.random((int(rng.integers(1, 300)), int(rng.integers(1, 300)))) < 0.15).astype(np.uint8) for _ in range(30)]
    got = ctx.sw_constrained(mats)
    exp = np.array([oracle.sw_constrained(B) for B in mats])
    assert np.array_equal(got.astype(np.float32), exp.astype(np.float32))


# ---- CRP: OTI, thresholds, binary plot bit-exact -------------------------------------------------
@pytest.mark.parametrize("m,tau", [(9, 1), (5, 2), (1, 1), (3, 3)])
def test_crp_bit_exact_small(ctx, oracle, m, tau):
    rng = np.random.default_rng(100 + m * 10 + tau)
    lens = [40, 37, 64, 129, 300, 31, 55, 270]
    tracks = _tracks(rng, lens)
    ctx.upload_tracks(tracks, m=m, tau=tau)
    prm = ctx.params(m=m, tau=tau)
    for (a, b) in [(0, 1), (1, 0), (2, 3), (4, 3), (5, 6), (7, 4), (3, 3)]:
        R, tq, tr, o = ctx.crp_pair(a, b, prm)
        Ro, info = oracle.chroma_cross_similarity(tracks[a], tracks[b], m=m, tau=tau, full=True)
        assert o == info["oti"]
        assert np.array_equal(tq, info["thr_q"]) and np.array_equal(tr, info["thr_r"])
        assert np.array_equal(R, Ro)


def test_crp_kappa_and_no_oti(ctx, oracle):
    rng = np.random.default_rng(77)
    tracks = _tracks(rng, [90, 120, 200])
    ctx.upload_tracks(tracks)
    for kappa in (0.0, 0.05, 0.1, 0.5, 1.0):
        for oti in (True, False):
            prm = ctx.params(kappa=kappa, oti=oti)
            R, tq, tr, o = ctx.crp_pair(0, 2, prm)
            Ro, info = oracle.chroma_cross_similarity(tracks[0], tracks[2], kappa=kappa, oti=oti, full=True)
            assert o == info["oti"] and np.array_equal(R, Ro)


def test_crp_exact_ties_duplicate_frames(ctx, oracle):
    """duplicated frames produce exactly equal distances: order statistics with ties"""
    rng = np.random.default_rng(9)
    a = rng.random((30, 12)).astype(np.float32)
    q = np.concatenate([a, a, a[:15]], 0)
    r = np.concatenate([a[5:], a[5:], np.zeros((12, 12), np.float32)], 0)
    ctx.upload_tracks([q, r])
    R, tq, tr, o = ctx.crp_pair(0, 1)
    Ro, info = oracle.chroma_cross_similarity(q, r, full=True)
    assert np.array_equal(tq, info["thr_q"]) and np.array_equal(tr, info["thr_r"]) and np.array_equal(R, Ro)


@pytest.mark.parametrize("mode", [1, 2, 3, 4, 5])
def test_crp_modes_bit_exact_medium(ctx, oracle, mode):
    """exact-everywhere strips (mode 1) and approximate + band + exact re-check (mode 2) give the
    same bits as the oracle; sizes large enough for the band path (NB >= 128), ragged tiles."""
    rng = np.random.default_rng(500 + mode)
    lens = [137, 200, 333, 498, 512, 523, 700, 1009, 150]
    tracks = _tracks(rng, lens, noise=0.1)
    ctx.upload_tracks(tracks)
    for kappa in (0.095, 0.3):
        prm = ctx.params(kappa=kappa, crp_mode=mode)
        for (a, b) in [(0, 1), (2, 3), (4, 5), (6, 7), (7, 2), (8, 6), (3, 3)]:
            R, tq, tr, o = ctx.crp_pair(a, b, prm)
            Ro, info = oracle.chroma_cross_similarity(tracks[a], tracks[b], kappa=kappa, full=True)
            assert o == info["oti"]
            assert np.array_equal(tq, info["thr_q"]) and np.array_equal(tr, info["thr_r"]), (mode, a, b)
            assert np.array_equal(R, Ro), (mode, a, b)


@pytest.mark.parametrize("opt", ["count_hh", "split_overlap"])
def test_crp_split_experimental_options_bit_exact(oracle, opt):
    """the phase-split pipeline's experimental switches (hi.hi-only count pass, selection on a second stream) change the schedule
    and the width of the candidate band, never the bits"""
    import acoss_b200
    rng = np.random.default_rng(91)
    lens = [300, 523, 700, 260]
    tracks = _tracks(rng, lens, noise=0.1)
    c = acoss_b200.Context(0)
    try:
        c.set_option(opt, 1)
        c.upload_tracks(tracks)
        prm = c.params(kappa=0.095, crp_mode=5)
        for (a, b) in [(0, 1), (1, 2), (2, 3)]:
            R, tq, tr, o = c.crp_pair(a, b, prm)
            Ro, info = oracle.chroma_cross_similarity(tracks[a], tracks[b], kappa=0.095, full=True)
            assert np.array_equal(tq, info["thr_q"]) and np.array_equal(tr, info["thr_r"]), (opt, a, b)
            assert np.array_equal(R, Ro), (opt, a, b)
    finally:
        c.close()


@pytest.mark.parametrize("cap", [0, 7, 100])
def test_crp_split_ambiguous_list_full(ctx, oracle, cap, monkeypatch):
    """phase-split pipeline: the batch's list of ambiguous cells (decided by crp_fix_kernel after the binarise pass) with room
    for `cap` cells only -- the strips that find it full decide their cells inside the pass; same bits as the oracle"""
    monkeypatch.setenv("ACOSS_AMB_CAP", str(cap))
    rng = np.random.default_rng(77)
    lens = [300, 523, 700, 260]
    tracks = _tracks(rng, lens, noise=0.1)
    ctx.upload_tracks(tracks)
    prm = ctx.params(kappa=0.095, crp_mode=5)
    for (a, b) in [(0, 1), (1, 2), (2, 3), (3, 0)]:
        R, tq, tr, o = ctx.crp_pair(a, b, prm)
        Ro, info = oracle.chroma_cross_similarity(tracks[a], tracks[b], kappa=0.095, full=True)
        assert np.array_equal(tq, info["thr_q"]) and np.array_equal(tr, info["thr_r"]), (cap, a, b)
        assert np.array_equal(R, Ro), (cap, a, b)


@pytest.mark.parametrize("mode", [4, 5])
@pytest.mark.parametrize("kappa", [0.004, 0.02, 0.5, 0.93, 1.0, 0.0])
def test_crp_gemm_extreme_percentiles(ctx, oracle, kappa, mode):
    """tensor-core kernel with brackets that are unbounded below / above (tiny and huge percentiles), ragged last tiles
    and strips, against the oracle"""
    rng = np.random.default_rng(77)
    tracks = _tracks(rng, [300, 521, 777], noise=0.12)
    ctx.upload_tracks(tracks)
    prm = ctx.params(kappa=kappa, crp_mode=mode)
    for (a, b) in [(0, 1), (1, 2), (2, 0)]:
        R, tq, tr, o = ctx.crp_pair(a, b, prm)
        Ro, info = oracle.chroma_cross_similarity(tracks[a], tracks[b], kappa=kappa, full=True)
        assert o == info["oti"]
        assert np.array_equal(tq, info["thr_q"]) and np.array_equal(tr, info["thr_r"]), (kappa, a, b)
        assert np.array_equal(R, Ro), (kappa, a, b)


def test_crp_band_degenerate_inputs(ctx, oracle):
    """constant / silent / duplicated frames: massive exact ties force the band kernel's fallbacks"""
    rng = np.random.default_rng(12)
    a = rng.random((40, 12)).astype(np.float32)
    const = np.tile(rng.random((1, 12)).astype(np.float32), (300, 1))
    silent = np.zeros((260, 12), np.float32)
    dup = np.concatenate([a] * 8, 0)
    half = np.concatenate([const[:150], rng.random((200, 12)).astype(np.float32)], 0)
    tracks = [const, silent, dup, half, _tracks(rng, [400])[0]]
    ctx.upload_tracks(tracks)
    for mode in (2, 3, 4, 5):
        prm = ctx.params(crp_mode=mode)
        for (x, y) in [(0, 0), (0, 1), (1, 1), (2, 2), (2, 4), (3, 0), (3, 2), (4, 3), (1, 4)]:
            R, tq, tr, o = ctx.crp_pair(x, y, prm)
            Ro, info = oracle.chroma_cross_similarity(tracks[x], tracks[y], full=True)
            assert o == info["oti"], (mode, x, y)
            assert np.array_equal(tq, info["thr_q"]) and np.array_equal(tr, info["thr_r"]), (mode, x, y)
            assert np.array_equal(R, Ro), (mode, x, y)


def test_band_vs_exact_mode_scores_large(ctx):
    """BASELINE-size tracks: mode 2 against mode 1 (itself checked against the oracle) on more pairs
    than the CPU oracle could do in seconds"""
    rng = np.random.default_rng(33)
    tracks = _tracks(rng, [int(x) for x in rng.integers(1600, 2400, size=10)], noise=0.08)
    ctx.upload_tracks(tracks)
    pairs = np.array([(i, j) for i in range(10) for j in range(10) if i < j], np.int32)
    q1, d1 = ctx.serra09_pairs(pairs, ctx.params(crp_mode=1), want_dmax=True)
    for mode in (2, 3, 4, 5):
        q2, d2 = ctx.serra09_pairs(pairs, ctx.params(crp_mode=mode), want_dmax=True)
        st = ctx.pair_stats()
        assert np.array_equal(q1, q2) and np.array_equal(d1, d2), mode
        # white-noise frames are the worst case for the spread-sample brackets (a missed bracket costs a whole exact row): < 4 %
        assert st["n_exact_cells"] < 0.04 * st["crp_cells"], (mode, st)
    R1 = ctx.crp_pair(0, 1, ctx.params(crp_mode=1))
    for mode in (2, 3, 4, 5):
        R2 = ctx.crp_pair(0, 1, ctx.params(crp_mode=mode))
        assert all(np.array_equal(x, y) for x, y in zip(R1[:3], R2[:3])), mode


def test_crp_golden(ctx, golden):
    g = golden("serra09_small.npz")
    n = int(g["n_tracks"])
    ctx.upload_tracks([g["ds%d" % i] for i in range(n)])
    for k, (a, b) in enumerate(g["crp_pairs"]):
        R, tq, tr, o = ctx.crp_pair(int(a), int(b))
        assert o == int(g["oti%d" % k])
        assert np.array_equal(tq, g["thrq%d" % k]) and np.array_equal(tr, g["thrr%d" % k])
        assert np.array_equal(R, g["crp%d" % k])


def test_too_short_track_errors(ctx):
    import acoss_b200
    ctx.upload_tracks([np.ones((9, 12), np.float32), np.ones((30, 12), np.float32)])
    with pytest.raises(acoss_b200.AcossError) as e:
        ctx.serra09_pairs(np.array([[0, 1]]))
    assert e.value.code == -3


# ---- pairs: scores ----------------------------------------------------------------------------
def test_pairs_scores_vs_oracle(ctx, oracle):
    rng = np.random.default_rng(31)
    lens = [int(x) for x in rng.integers(30, 420, size=14)]
    tracks = _tracks(rng, lens)
    ctx.upload_tracks(tracks)
    pairs = np.array([(i, j) for i in range(14) for j in range(14) if i != j], np.int32)
    q, d = ctx.serra09_pairs(pairs, want_dmax=True)
    oq, od, cells = oracle.serra09_pairs(tracks, pairs, want_dmax=True, n_threads=8)
    assert np.array_equal(q, oq) and np.array_equal(d, od)
    st = ctx.pair_stats()
    assert st["dp_cells"] == 2 * cells and st["n_launches"] >= 5


def test_pairs_scores_long_tracks_sample(ctx, oracle):
    """BASELINE config sizes (n ~ 2000 at the DP stage): a few pairs against the oracle."""
    rng = np.random.default_rng(32)
    tracks = _tracks(rng, [2400, 1600, 2011, 1893])
    ctx.upload_tracks(tracks)
    pairs = np.array([(0, 1), (1, 0), (2, 3), (0, 3)], np.int32)
    q, d = ctx.serra09_pairs(pairs, want_dmax=True)
    oq, od, _ = oracle.serra09_pairs(tracks, pairs, want_dmax=True, n_threads=4)
    assert np.array_equal(q, oq) and np.array_equal(d, od)


def test_pairs_multibatch_and_alloc_retry_vs_oracle(oracle):
    """The multi-batch path of the pair pipeline (scratch-buffer reuse and regrowth across batches, per-batch DP
    hand-out counters, kernel choice per batch on ragged lengths straddling 512 frames) and the halve-the-budget
    retry after an allocation failure: scores bit-equal to the single-batch run and to the oracle."""
    import acoss_b200
    rng = np.random.default_rng(77)
    lens = [int(x) for x in rng.integers(60, 500, size=8)] + [int(x) for x in rng.integers(520, 760, size=6)]
    tracks = _tracks(rng, lens)
    n = len(tracks)
    pairs = np.array([(i, j) for i in range(n) for j in range(n) if i != j], np.int32)
    pairs = pairs[rng.permutation(len(pairs))[:90]]
    oq, od, _ = oracle.serra09_pairs(tracks, pairs, want_dmax=True, n_threads=8)
    c = acoss_b200.Context(0)
    try:
        c.upload_tracks(tracks)
        q1, d1 = c.serra09_pairs(pairs, want_dmax=True)
        assert c.pair_stats()["n_batches"] == 1
        c.set_option("crp_budget_bytes", 1 << 20)          # ~2-20 pairs per batch
        q2, d2 = c.serra09_pairs(pairs, want_dmax=True)
        st = c.pair_stats()
        assert st["n_batches"] > 5, st
        assert "gemm" in st["crp_kernels"] or "split" in st["crp_kernels"], st
        short = np.array([(i, j) for i in range(8) for j in range(8) if i != j], np.int32)     # tracks below the 512-frame switch
        q4, d4 = c.serra09_pairs(short, want_dmax=True)
        st = c.pair_stats()
        assert st["n_batches"] > 3 and st["crp_kernels"] == ["band"], st
        sq, sd_, _ = oracle.serra09_pairs(tracks, short, want_dmax=True, n_threads=8)
        assert np.array_equal(q4, sq) and np.array_equal(d4, sd_)
        assert np.array_equal(q1, oq) and np.array_equal(d1, od)
        assert np.array_equal(q2, oq) and np.array_equal(d2, od)
        c.set_option("crp_budget_bytes", 64 << 20)
        c.set_option("inject_alloc_failures", 2)            # two failed reservations -> budget quartered, call succeeds
        q3, d3 = c.serra09_pairs(pairs, want_dmax=True)
        assert c.pair_stats()["n_batches"] >= 2
        assert np.array_equal(q3, oq) and np.array_equal(d3, od)
        with pytest.raises(acoss_b200.AcossError):
            c.set_option("no_such_option", 1)
    finally:
        c.close()


def test_median_downsample_vs_oracle(ctx, oracle):
    rng = np.random.default_rng(41)
    raws = [rng.random((n, 12)).astype(np.float32) for n in (1, 39, 40, 41, 1600, 2399, 80)]
    for fac in (40, 7, 2):
        got = ctx.median_downsample(raws, fac)
        for x, y in zip(got, raws):
            assert np.array_equal(x, oracle.sync_median(y, fac))


# ---- N x N stage + the reference-facing classes ------------------------------------------------
def _write_golden_dataset(g, workdir):
    from acoss_b200 import synth
    n = int(g["n_tracks"])
    labels = [str(x) for x in g["labels"]]
    ds = dict(labels=labels, work_ids=labels, track_ids=["P_%d" % i for i in range(n)],
              hpcp=[g["hpcp%d" % i] for i in range(n)], fac=int(g["fac"]))
    fdir = synth.write_dataset(str(workdir / "feat"), ds, str(workdir / "ds.csv"))
    return str(workdir / "ds.csv"), fdir


def test_serra09_class_matches_reference_golden(golden, workdir):
    """The reference's own Serra09 class (run verbatim on the essentia restatement, frozen in
    tests/golden) vs the drop-in class: Ds before / after normalisation bit-equal, eval identical."""
    import acoss_b200
    g = golden("serra09_small.npz")
    csv, fdir = _write_golden_dataset(g, workdir)
    s = acoss_b200.Serra09(csv, fdir, shortname="g")
    for i in range(s.N):
        assert np.array_equal(s.load_features(i), g["ds%d" % i])
    s.all_pairwise(1, n_cores=4, symmetric=True)
    assert np.array_equal(np.array(s.Ds["main"]), g["serra_raw"])
    s.normalize_by_length()
    assert np.array_equal(np.array(s.Ds["main"]), g["serra_norm"])
    MR, MRR, MDR, MAP, tops = s.getEvalStatistics("main")
    assert [MR, MRR, MDR, MAP] == g["serra_eval"].tolist()
    s.cleanup_memmap()


def test_serra09_L48_full_run_vs_golden(golden, workdir):
    """A full Serra09 run at the L length regime (48 tracks, 1128 pairs, M, N ~ 1600-2400, the HARD cover model so
    that MAP < 1): raw and normalised matrices bit-equal to the oracle's / the reference class's, eval tuple
    identical (golden: oracle/make_golden.py gen_serra_L48)."""
    import acoss_b200
    from acoss_b200 import synth
    from oracle.make_golden import serra_L48_dataset
    g = golden("serra09_L48.npz")
    ds = serra_L48_dataset()
    assert np.array_equal([float(np.sum(h, dtype=np.float64)) for h in ds["hpcp"]], g["checksum"])
    fdir = synth.write_dataset(str(workdir / "feat"), ds, str(workdir / "ds.csv"))
    s = acoss_b200.Serra09(str(workdir / "ds.csv"), fdir, shortname="gL", downsample_fac=1)
    s.all_pairwise(1, symmetric=True)
    assert "gemm" in s.last_stats["crp_kernels"] or "split" in s.last_stats["crp_kernels"]
    assert np.array_equal(np.array(s.Ds["main"]), g["serra_raw"])
    s.normalize_by_length()
    assert np.array_equal(np.array(s.Ds["main"]), g["serra_norm"])
    MR, MRR, MDR, MAP, tops = s.getEvalStatistics("main")
    assert [MR, MRR, MDR, MAP] == g["serra_eval"].tolist()
    assert MAP < 0.95 and MR > 1.0          # the check is not vacuous: retrieval is not saturated
    # resume path: all_pairwise(precomputed=True) re-loads `<prefix>_Ds.h5` (algorithm_template.py:163-166)
    s2 = acoss_b200.Serra09(str(workdir / "ds.csv"), fdir, shortname="gL", downsample_fac=1)
    assert not np.any(np.array(s2.Ds["main"]))              # the constructor truncates the memmap
    s2.all_pairwise(1, symmetric=True, precomputed=True)
    assert np.array_equal(np.array(s2.Ds["main"]), g["serra_raw"])
    assert sum(len(c) for c in s2.cliques.values()) == s2.N
    s2.Ds["main"] = np.array(s2.Ds["main"])
    s2.preload()
    s2.normalize_by_length()
    assert [float(x) for x in s2.getEvalStatistics("main")[:4]] == g["serra_eval"].tolist()
    ctx2 = s.context()
    q, d = ctx2.serra09_pairs(g["pairs"][:64], want_dmax=True)
    assert np.array_equal(d, g["dmax"][:64]) and np.array_equal(q, g["qmax"][:64])
    s.cleanup_memmap()


def test_eval_vs_reference_golden(ctx, golden):
    import torch
    g = golden("eval_reference.npz")
    for c in range(int(g["n"])):
        if int(g["ties%d" % c]):
            continue
        D, labels, res = g["D%d" % c], [str(x) for x in g["labels%d" % c]], g["res%d" % c]
        cl = {}
        for i, l in enumerate(labels):
            cl.setdefault(l, set([])).add(i)
        # list(set) like the reference (algorithm_template.py:220): the within-clique order fixes the
        # order in which nanmean sums AllMap, hence the last ulp of MAP
        cliques = [list(c) for c in cl.values()]
        Ks = np.array([len(x) for x in cliques])
        idx = np.argsort(-Ks)
        cliques = [cliques[i] for i in idx]
        Ks = Ks[idx]
        order = np.concatenate(cliques).astype(np.int32)
        csize = np.repeat(Ks, Ks).astype(np.int32)
        cstart = np.repeat(np.concatenate([[0], np.cumsum(Ks)[:-1]]), Ks).astype(np.int32)
        Dd = torch.from_numpy(D).cuda()
        torch.cuda.synchronize()
        out, ranks = ctx.eval_statistics(Dd.data_ptr(), D.shape[0], order, csize, cstart, [1, 10, 100, 1000])
        assert out.tolist() == res.tolist()


def test_eval_vs_oracle_with_ties_stable(ctx, oracle):
    import torch
    rng = np.random.default_rng(3)
    n = 90
    labels = ["W%d" % (i // 3) for i in range(60)] + ["S%d" % i for i in range(30)]
    D = (np.round(rng.random((n, n)) * 6) / 6).astype(np.float32)
    cl = {}
    for i, l in enumerate(labels):
        cl.setdefault(l, []).append(i)
    exp = oracle.eval_statistics(D, cl, stable=True)
    cliques = list(cl.values())
    Ks = np.array([len(x) for x in cliques])
    idx = np.argsort(-Ks, kind="stable")
    cliques = [cliques[i] for i in idx]
    Ks = Ks[idx]
    order = np.concatenate(cliques).astype(np.int32)
    csize = np.repeat(Ks, Ks).astype(np.int32)
    cstart = np.repeat(np.concatenate([[0], np.cumsum(Ks)[:-1]]), Ks).astype(np.int32)
    Dd = torch.from_numpy(D).cuda()
    torch.cuda.synchronize()
    out, _ = ctx.eval_statistics(Dd.data_ptr(), n, order, csize, cstart, [1, 10, 100, 1000])
    assert out[:4].tolist() == list(exp[:4]) and out[4:].tolist() == exp[4].tolist()


def test_benchmark_entry_point(golden, workdir):
    import acoss_b200
    g = golden("serra09_small.npz")
    csv, fdir = _write_golden_dataset(g, workdir)
    res = acoss_b200.benchmark(csv, fdir, algorithm="Serra09", shortname="bm")
    assert list(res["main"][:4]) == g["serra_eval"].tolist()
    assert (workdir / "results_bm_Serra09.csv").exists()


# ---- similarity network fusion + LateFusionChen ------------------------------------------------
@pytest.mark.gpu
def test_snf_vs_reference_golden(ctx, golden):
    """doSimilarityFusion run verbatim from the reference (float32 inputs, incl. the Q8 aliasing)."""
    g = golden("snf_reference.npz")
    for c in range(int(g["n"])):
        nm, N, K, nit = (int(x) for x in g["cfg%d" % c])
        mats = [g["D%d_%d" % (c, k)] for k in range(nm)]
        Ws, fused = ctx.snf(mats, K=K, niters=nit, reg_diag=1)
        assert fused.dtype == np.float64 and fused.shape == (N, N)
        # tolerance: the affinity stage is float32 (expf vs numpy's exp differ in the last ulp)
        np.testing.assert_allclose(fused, g["fused%d" % c], rtol=2e-5, atol=1e-9)


@pytest.mark.gpu
def test_snf_float64_vs_numpy_restatement(ctx):
    """float64 inputs: compare with a plain numpy restatement of the same equations written here
    (Jacobi first sweep, Gauss-Seidel afterwards)."""
    rng = np.random.default_rng(11)
    N, K, nit, nm = 150, 10, 6, 3
    mats = []
    for _ in range(nm):
        X = rng.random((N, 5))
        mats.append(np.sqrt(((X[:, None] - X[None]) ** 2).sum(-1)) + 0.05 * rng.random((N, N)))
    Ws_dev, fused = ctx.snf(mats, K=K, niters=nit, reg_diag=1)

    def getW(D):
        S = 0.5 * (D + D.T)
        np.fill_diagonal(S, 0)
        md = np.sort(S, 1)[:, :K + 1].mean(1) * (K + 1) / K
        eps = (md[:, None] + md[None, :] + S) / 3
        den = 2 * (0.5 * eps) ** 2
        den[den == 0] = 1
        return np.exp(-S ** 2 / den)
    Ws = [getW(D) for D in mats]
    for a, b in zip(Ws, Ws_dev):
        np.testing.assert_allclose(b, a, rtol=1e-12)
    Ps = [W / W.sum(1, keepdims=True) for W in Ws]
    Ss = []
    for W in Ws:
        S = np.zeros_like(W)
        J = np.argsort(-W, 1, kind="stable")[:, :K]
        for i in range(N):
            S[i, J[i]] = W[i, J[i]] / W[i, J[i]].sum()
        Ss.append(S)
    Pts = [P.copy() for P in Ps]
    for it in range(nit):
        src = [P.copy() for P in Pts] if it == 0 else Pts
        for i in range(nm):
            X = sum(src[k] for k in range(nm) if k != i) / (nm - 1)
            Pts[i] = Ss[i] @ X @ Ss[i].T + np.eye(N)
            if it > 0:
                src = Pts
    np.testing.assert_allclose(fused, sum(Pts) / nm, rtol=1e-10)


@pytest.mark.gpu
def test_snf_argument_errors(ctx):
    import acoss_b200
    D = np.random.default_rng(0).random((10, 10)).astype(np.float32)
    with pytest.raises(acoss_b200.AcossError):
        ctx.snf([D], K=3)                      # one matrix: the reference divides by N-1 = 0
    with pytest.raises(acoss_b200.AcossError):
        ctx.snf([D, D], K=9)                   # kth = K+1 out of bounds (np.partition raises)


@pytest.mark.gpu
def test_chen_class_matches_reference_golden(golden, workdir):
    """The reference's ChenFusion run verbatim (on the essentia restatement) vs the drop-in class."""
    import acoss_b200
    g = golden("serra09_small.npz")
    csv, fdir = _write_golden_dataset(g, workdir)
    c = acoss_b200.ChenFusion(csv, fdir, shortname="g")
    c.all_pairwise(1, n_cores=4, symmetric=True)
    assert np.array_equal(np.array(c.Ds["qmax"]), g["chen_q_raw"])
    assert np.array_equal(np.array(c.Ds["dmax"]), g["chen_d_raw"])
    c.normalize_by_length()
    assert np.array_equal(np.array(c.Ds["qmax"]), g["chen_q_norm"])
    assert np.array_equal(np.array(c.Ds["dmax"]), g["chen_d_norm"])
    c.do_late_fusion()
    np.testing.assert_allclose(np.array(c.Ds["Late"]), g["chen_late"], rtol=2e-5, atol=1e-9)
    assert np.array_equal(np.array(c.Ds["qmax"]), -g["chen_q_norm"])
    for k in ("qmax", "dmax", "Late"):
        res = c.getEvalStatistics(k)
        assert list(res[:4]) == g["chen_eval_" + k].tolist()
    c.cleanup_memmap()


@pytest.mark.gpu
def test_benchmark_entry_point_chen(golden, workdir):
    import acoss_b200
    g = golden("serra09_small.npz")
    csv, fdir = _write_golden_dataset(g, workdir)
    res = acoss_b200.benchmark(csv, fdir, algorithm="LateFusionChen", shortname="bm")
    for k in ("qmax", "dmax", "Late"):
        assert list(res[k][:4]) == g["chen_eval_" + k].tolist()


# ---- SiMPle and FTM2D ------------------------------------------------------------------------
def _simple_ftm_dataset(workdir):
    from acoss_b200 import synth
    from oracle.make_golden import simple_ftm_dataset
    ds, extra = simple_ftm_dataset()
    fdir = synth.write_dataset(str(workdir / "feat"), ds, str(workdir / "ds.csv"), extra=extra)
    return str(workdir / "ds.csv"), fdir


def test_serra09_class_from_packed_store_identical(workdir):
    """Serra09 through the packed feature store (one mmap for the whole dataset, featio.save_packed) gives the same
    matrix, bit for bit, as through the per-track files, including the clique bookkeeping and the eval tuple."""
    import acoss_b200
    from acoss_b200 import synth
    ds = synth.make_named("covers80", "S", seed=11, scale=0.1)
    fdir = synth.write_dataset(str(workdir / "feat"), ds, str(workdir / "ds.csv"))
    a = acoss_b200.Serra09(str(workdir / "ds.csv"), fdir, shortname="pk_a", downsample_fac=ds["fac"])
    a.all_pairwise(1, symmetric=True)
    a.normalize_by_length()
    ra = a.getEvalStatistics("main")
    b = acoss_b200.Serra09(str(workdir / "ds.csv"), fdir, shortname="pk_b", downsample_fac=ds["fac"])
    path = b.build_packed_store()
    assert os.path.exists(path) and path == b.packed_store_path()
    c = acoss_b200.Serra09(str(workdir / "ds.csv"), fdir, shortname="pk_b", downsample_fac=ds["fac"])
    c.filepaths = ["/nonexistent/%d.h5" % k for k in range(c.N)]        # the per-track files must not be touched
    c.all_pairwise(1, symmetric=True)
    c.normalize_by_length()
    rc = c.getEvalStatistics("main")
    assert np.array_equal(np.array(a.Ds["main"]), np.array(c.Ds["main"]))
    assert {k: sorted(v) for k, v in a.cliques.items()} == {k: sorted(v) for k, v in c.cliques.items()}
    assert list(ra[:4]) == list(rc[:4])
    for o in (a, b, c):
        o.cleanup_memmap()



def test_simple_pairs_vs_reference_golden(ctx, golden):
    """-simple_sim(Si, oti(Si, Sj)) of the reference (float64) vs the matrix-profile kernel.
    Tolerance: the reference accumulates FFT-seeded sliding dot products (~1e-13 absolute noise)."""
    g = golden("simple_ftm2d_reference.npz")
    feats = [g["simple_feat%d" % i] for i in range(10)]
    ctx.simple_upload(feats)
    got = ctx.simple_pairs(g["simple_pairs"])
    np.testing.assert_allclose(got, g["simple_scores"], rtol=1e-9, atol=1e-11)
    n = len(feats)
    allp = np.array([(i, j) for i in range(n) for j in range(n) if i != j], np.int32)
    sc = ctx.simple_pairs(allp)
    D = np.zeros((n, n), np.float32)
    D[allp[:, 0], allp[:, 1]] = sc
    np.testing.assert_allclose(D, g["simple_D"], rtol=1e-6, atol=1e-9)


def test_simple_pairs_large_vs_reference_golden(ctx, golden):
    """BASELINE C5 size: 8 tracks of 170-246 columns, all 56 ordered pairs against the reference's simple_sim run
    verbatim (float64; tolerance as above)."""
    g = golden("simple_large.npz")
    feats = [g["feat%d" % i] for i in range(int(g["n"]))]
    assert min(f.shape[1] for f in feats) >= 150 and max(f.shape[1] for f in feats) <= 250
    ctx.simple_upload(feats)
    got = ctx.simple_pairs(g["pairs"])
    np.testing.assert_allclose(got, g["scores"], rtol=1e-9, atol=1e-11)


def test_simple_preprocess_vs_reference_golden(ctx, golden):
    """Device WIN/SKIP mean + Hann smoothing + column normalisation vs the reference's Simple.load_features run
    verbatim (simple_silva.py:34-43, 56-66; golden made by oracle/make_golden.py).  float64: tolerance 1e-12 (the
    summation order inside numpy.convolve is not specified); plus ragged / short inputs against the host mirror."""
    from oracle.make_golden import simple_ftm_dataset
    import acoss_b200
    g = golden("simple_ftm2d_reference.npz")
    ds, _ = simple_ftm_dataset()
    got = ctx.simple_preprocess(ds["hpcp"])
    assert len(got) == 10
    for i, f in enumerate(got):
        ref = g["simple_feat%d" % i]
        assert f.shape == ref.shape
        np.testing.assert_allclose(f, ref, rtol=1e-12, atol=1e-15)
    # ragged: T not a multiple of SKIP, last window clipped, a track shorter than SKIP (no columns), other WIN / SKIP
    from oracle import ports as P
    rng = np.random.default_rng(5)
    raw = [rng.random((n, 12)).astype(np.float32) for n in (1234, 99, 100, 777, 2001)]
    for win, skip in ((200, 100), (64, 50), (30, 70)):
        got = ctx.simple_preprocess(raw, win=win, skip=skip)
        for r, f in zip(raw, got):
            exp = P.simple_features(r.T, WIN=win, SKIP=skip)
            assert f.shape == exp.shape
            if exp.shape[1]:
                np.testing.assert_allclose(f, exp, rtol=1e-12, atol=1e-15)


def test_simple_generic_sslen_and_errors(ctx):
    import acoss_b200
    rng = np.random.default_rng(5)
    feats = []
    for L in (31, 24, 40):
        x = rng.random((12, L)) + 0.05
        feats.append(x / np.sqrt((x ** 2).sum(0, keepdims=True)))
    ctx.simple_upload(feats)
    pairs = np.array([(0, 1), (1, 2), (2, 0)], np.int32)

    def brute(a, b, ss):
        prof_a, prof_b = a.sum(1), b.sum(1)
        v = [np.dot(prof_a, np.roll(prof_b, k)) for k in range(12)]
        b = np.roll(b, int(np.argsort(v)[-1]), axis=0)
        na, nb = a.shape[1] - ss + 1, b.shape[1] - ss + 1
        mp = [min(((a[:, p:p + ss] - b[:, q:q + ss]) ** 2).sum() for q in range(nb)) for p in range(na)]
        return -np.median(mp)
    for ss in (10, 7):
        got = ctx.simple_pairs(pairs, sslen=ss)
        exp = [brute(feats[a], feats[b], ss) for a, b in pairs]
        np.testing.assert_allclose(got, exp, rtol=1e-12, atol=1e-14)
    with pytest.raises(acoss_b200.AcossError):
        ctx.simple_pairs(pairs, sslen=30)          # track 1 has 24 columns < SSLEN


def test_ftm2d_shingles_vs_reference_golden(ctx, golden, workdir):
    """Per-track FTM2D shingles on the device (beat-synchronous median, chrompwr, 12 x 75 2-D DFT magnitudes, log, median, unit
    norm) against the reference's FTM2D.load_features run verbatim.  Tolerance 2e-5 of the largest entry: the reference's
    scipy.fftpack.fft2 of the float32 chroma is complex64 (its magnitudes carry float32 FFT rounding), the device accumulates
    the same DFT sums in float64 and rounds the magnitude to float32 once.  Also through the classes (cache files,
    Simple.load_features / smooth on the device) and the error for a track with fewer beats than WIN."""
    import acoss_b200
    from oracle.make_golden import simple_ftm_dataset
    from oracle import ports as P
    g = golden("simple_ftm2d_reference.npz")
    ds, extra = simple_ftm_dataset()
    ons = [m["onsets"] for m in extra["madmom_features"]]
    S = ctx.ftm2d_shingles(ds["hpcp"], ons)
    assert S.shape == (10, 900)
    for i in range(10):
        ref = g["ftm_shingle%d" % i]
        np.testing.assert_allclose(S[i], ref, rtol=0, atol=2e-5 * np.max(np.abs(ref)))
        assert abs(np.sqrt(np.sum(S[i] ** 2)) - 1.0) < 1e-12
    with pytest.raises(acoss_b200.AcossError) as e:
        ctx.ftm2d_shingles([ds["hpcp"][0][:600]], [ons[0][ons[0] < 600]])
    assert e.value.code == -3
    csv, fdir = _simple_ftm_dataset(workdir)
    f = acoss_b200.FTM2D(csv, fdir, shortname="gs")
    np.testing.assert_allclose(f.load_features(3), g["ftm_shingle3"], rtol=0, atol=2e-5 * np.max(np.abs(g["ftm_shingle3"])))
    f2 = acoss_b200.FTM2D(csv, fdir, shortname="gs")           # second object: served from the cache file of the first
    assert np.array_equal(f2.load_features(3), f.load_features(3)) and len(f2.cliques) == 1
    s = acoss_b200.Simple(csv, fdir, shortname="gs")
    for i in (0, 7):
        np.testing.assert_allclose(s.load_features(i), g["simple_feat%d" % i], rtol=1e-12, atol=1e-15)
    x = np.random.default_rng(1).random((12, 40)).astype(np.float32).astype(np.float64)
    np.testing.assert_allclose(s.smooth(x), P.simple_features(x, WIN=1, SKIP=1), rtol=1e-12, atol=1e-15)


def test_ftm2d_allpairs_vs_reference_golden(ctx, golden):
    g = golden("simple_ftm2d_reference.npz")
    S = np.stack([g["ftm_shingle%d" % i] for i in range(10)])
    D = ctx.ftm2d_allpairs(S)
    exp = np.exp(-((S[:, None, :] - S[None, :, :]) ** 2).sum(-1))
    np.testing.assert_allclose(D, exp, rtol=1e-13)
    ref = g["ftm_D"]                                # float32 memmap of the reference, diagonal 0
    Dz = D.astype(np.float32)
    np.fill_diagonal(Dz, 0)
    assert np.array_equal(Dz, ref)
    S2 = np.random.default_rng(2).random((131, 900))
    D2 = ctx.ftm2d_allpairs(S2)
    np.testing.assert_allclose(D2, np.exp(-((S2 ** 2).sum(1)[:, None] + (S2 ** 2).sum(1)[None] - 2 * S2 @ S2.T)), rtol=1e-9)


def test_simple_and_ftm2d_classes_and_benchmark(golden, workdir):
    import acoss_b200
    g = golden("simple_ftm2d_reference.npz")
    csv, fdir = _simple_ftm_dataset(workdir)
    res = acoss_b200.benchmark(csv, fdir, algorithm="SiMPle", shortname="g")
    assert list(res["main"][:4]) == g["simple_eval"].tolist()
    res = acoss_b200.benchmark(csv, fdir, algorithm="FTM2D", shortname="g")
    assert list(res["main"][:4]) == g["ftm_eval"].tolist()
    f = acoss_b200.FTM2D(csv, fdir, shortname="g2")
    f.all_pairwise(1, symmetric=True)
    # shingles now come from the device (float32-FFT-rounding tolerance, see test_ftm2d_shingles_vs_reference_golden)
    np.testing.assert_allclose(np.array(f.Ds["main"]), g["ftm_D"], rtol=2e-5, atol=1e-7)
    f.cleanup_memmap()


# ---- EarlyFusionTralie -----------------------------------------------------------------------
def _early_dataset(workdir):
    from acoss_b200 import synth
    from oracle.make_golden import early_dataset
    ds = early_dataset()
    fdir = synth.write_dataset(str(workdir / "feat"), ds, str(workdir / "ds.csv"))
    return str(workdir / "ds.csv"), fdir, ds


def test_earlyfusion_pair_matrices_vs_reference_golden(ctx, golden):
    """CSMs / early matrix (float32) and the four binary matrices of two pairs against the reference's
    free functions run verbatim.  The GEMM summation order differs from numpy's BLAS, so the float
    matrices agree to float32 rounding; the binary matrices must then be identical (seeded data has no
    near-ties at the k-th neighbour)."""
    from oracle.make_golden import early_dataset
    g = golden("earlyfusion_reference.npz")
    ds = early_dataset()
    assert np.array_equal([float(np.sum(f["mfccs"], dtype=np.float64) + np.sum(f["chromas"], dtype=np.float64))
                           for f in ds["block_feats"]], g["checksum"])
    ctx.earlyfusion_upload(ds["block_feats"])
    for k, (a, b) in enumerate(g["mat_pairs"]):
        csm, bn, sc = ctx.earlyfusion_pair_matrices(int(a), int(b), kappa=0.1, K=10)
        np.testing.assert_allclose(csm[:3], g["csm%d" % k][:3], rtol=2e-5, atol=2e-6)
        np.testing.assert_allclose(csm[3], g["csm%d" % k][3], rtol=1e-4, atol=1e-6)
        assert np.array_equal(bn.sum(2), g["bin%d" % k].sum(2))          # exactly k ones per row
        mism = int((bn != g["bin%d" % k]).sum())
        assert mism == 0, "binary matrices differ in %d cells" % mism
        assert sc.astype(np.float32).tolist() == [g["D_" + s][a, b] for s in ("mfccs", "ssms", "chromas", "early")]


def test_earlyfusion_large_vs_reference_golden(ctx, golden):
    """BASELINE C3 size (367-587 blocks, d = 1000 / 1225 / 480, k = 37-59 neighbours): CSM rows, the four binary
    matrices and the four scores of four pairs against the reference's functions run verbatim.

    The reference's CSMs come out of numpy's BLAS sgemm; no other implementation reproduces its summation order, so
    a cell whose value is within float32-GEMM rounding of its row's k-th neighbour can legitimately fall on either
    side.  The golden stores every row's k-th and (k+1)-th smallest value (float64 sort of the reference's matrix):
    a differing cell is accepted only in a row whose gap is below the rounding tolerance, and then the scores are
    not compared for that matrix.  Everything else must be identical."""
    from oracle.make_golden import early_large_dataset
    g = golden("earlyfusion_large.npz")
    ds = early_large_dataset()
    bf = ds["block_feats"]
    assert np.array_equal([float(np.sum(f["mfccs"], dtype=np.float64) + np.sum(f["ssms"], dtype=np.float64) +
                                 np.sum(f["chromas"], dtype=np.float64)) for f in bf], g["checksum"])
    ctx.earlyfusion_upload(bf)
    n_diff_total = 0
    for k, (a, b) in enumerate(g["mat_pairs"]):
        M, N, kk = (int(x) for x in g["shape%d" % k])
        csm, bn, sc = ctx.earlyfusion_pair_matrices(int(a), int(b), kappa=0.1, K=10)
        assert csm.shape == (4, M, N)
        np.testing.assert_allclose(csm[:3, ::16], g["csm_rows%d" % k][:3], rtol=2e-5, atol=2e-6)
        np.testing.assert_allclose(csm[3, ::16], g["csm_rows%d" % k][3], rtol=2e-4, atol=1e-6)
        ref_bin = np.unpackbits(g["binpk%d" % k], axis=2)[:, :, :N]
        assert np.array_equal(bn.sum(2), ref_bin.sum(2))                  # exactly k ones per row
        for t in range(4):
            rows = np.nonzero((bn[t] != ref_bin[t]).any(1))[0]
            n_diff_total += len(rows)
            kth = g["kth%d" % k][t]
            for r in rows:          # only a near-tie at the k-th neighbour may differ
                gap = kth[r, 1] - kth[r, 0]
                tol = (2e-5 if t < 3 else 2e-4) * max(abs(kth[r, 1]), 1e-3)
                assert gap <= tol, "pair %d matrix %d row %d differs with a gap of %g" % (k, t, r, gap)
                assert int((bn[t, r] != ref_bin[t, r]).sum()) == 2
            if len(rows) == 0:
                assert np.float32(sc[t]) == np.float32(g["scores%d" % k][t]), (k, t, sc[t], g["scores%d" % k][t])
                assert np.float32(sc[t]) == g[("D_mfccs", "D_ssms", "D_chromas", "D_early")[t]][a, b]
    # whole-batch call (the production path: batched kernels instead of the test hook)
    prs = np.array([(int(a), int(b)) for a, b in g["mat_pairs"]], np.int32)
    got = ctx.earlyfusion_pairs(prs, kappa=0.1, K=10)
    for k in range(len(prs)):
        _, _, sc = ctx.earlyfusion_pair_matrices(int(prs[k, 0]), int(prs[k, 1]), kappa=0.1, K=10)
        assert got[k].tolist() == sc.tolist()
    print("earlyfusion_large: rows with a near-tie difference: %d" % n_diff_total)
    assert n_diff_total <= 8


def _raw_early_tracks(seed, n_tracks):
    """raw inputs of EarlyFusion.load_features: mfcc_htk (20, T) with a few NaNs, chroma (T, 12), beat onsets"""
    rng = np.random.default_rng(seed)
    out = []
    for t in range(n_tracks):
        T = int(rng.integers(1500, 2200))
        mf = (np.cumsum(rng.standard_normal((20, T)), 1) * 0.1 + rng.standard_normal((20, T))).astype(np.float32)
        mf[rng.integers(0, 20), rng.integers(0, T)] = np.nan
        ch = rng.random((T, 12)).astype(np.float32)
        gaps = rng.integers(18, 45, size=T // 18 + 2)            # 18-44 frames per beat: blocks of ~ 360-880 frames -> 50 / 40
        on = np.cumsum(gaps)
        out.append((mf, ch, on[on < T - 1].astype(np.int64)))
    return out


def test_earlyfusion_block_features_vs_port(ctx, workdir):
    """EarlyFusion.load_features on the device (blocks, anti-aliased resize, z-normalisation, SSM triangle, chroma blocks, chroma
    median) against oracle/ports.py:early_block_features -- a restatement whose resize uses the real scipy.ndimage Gaussian
    filter and restates skimage's order-1 warp (skimage is absent: this stage's parity is UNPINNED).  float32 outputs of float64
    arithmetic: rtol 2e-6 (summation order of the 20- / 50-term sums), SSM entries near 0 to 2e-6 absolute.  Then the class:
    benchmark(\"EarlyFusionTraile\") runs from a RAW feature directory."""
    import acoss_b200
    from acoss_b200 import synth
    from oracle import ports as P
    raw = _raw_early_tracks(11, 3)
    got = ctx.earlyfusion_block_features([m.T for m, _, _ in raw], [c for _, c, _ in raw], [o for _, _, o in raw])
    for (mf, ch, on), g in zip(raw, got):
        exp = P.early_block_features(mf, ch, on)
        assert g["mfccs"].shape == exp["mfccs"].shape == (len(on) - 20, 1000)
        np.testing.assert_allclose(g["mfccs"], exp["mfccs"], rtol=2e-6, atol=1e-7)
        np.testing.assert_allclose(g["ssms"], exp["ssms"], rtol=2e-6, atol=2e-6)
        np.testing.assert_allclose(g["chromas"], exp["chromas"], rtol=2e-6, atol=1e-7)
        assert np.array_equal(g["chroma_med"], exp["chroma_med"])
    # upsampling blocks (fewer frames than output rows: no anti-alias filter, zero-padded borders)
    short = [(raw[0][0][:, :400], raw[0][1][:400], np.arange(0, 400, 2).astype(np.int64))]
    g2 = ctx.earlyfusion_block_features([short[0][0].T], [short[0][1]], [short[0][2]])[0]
    e2 = P.early_block_features(*short[0])
    np.testing.assert_allclose(g2["mfccs"], e2["mfccs"], rtol=2e-6, atol=1e-7)
    np.testing.assert_allclose(g2["chromas"], e2["chromas"], rtol=2e-6, atol=1e-7)
    with pytest.raises(acoss_b200.AcossError) as e:
        ctx.earlyfusion_block_features([raw[0][0].T], [raw[0][1]], [raw[0][2][:15]])
    assert e.value.code == -3
    # the entry point from raw feature files (12 tracks: K + 1 = 21 neighbours need more than 21 songs for the fusion -> 24)
    raw = _raw_early_tracks(12, 24)
    labels = ["W_%d" % (k // 2) for k in range(24)]
    ds = dict(labels=labels, work_ids=labels, track_ids=["P_%d" % k for k in range(24)], hpcp=[c for _, c, _ in raw], fac=1)
    extra = {"mfcc_htk": [m for m, _, _ in raw], "madmom_features": [{"onsets": o} for _, _, o in raw]}
    fdir = synth.write_dataset(str(workdir / "feat"), ds, str(workdir / "ds.csv"), extra=extra)
    res = acoss_b200.benchmark(str(workdir / "ds.csv"), fdir, algorithm="EarlyFusionTraile", shortname="raw")
    assert set(res) == {"mfccs", "ssms", "chromas", "early", "late", "early+late"}
    e = acoss_b200.EarlyFusion(str(workdir / "ds.csv"), fdir, shortname="raw")
    f5 = e.load_features(5)                                    # served from the cache file written by the run above
    np.testing.assert_allclose(f5["mfccs"], P.early_block_features(*raw[5])["mfccs"], rtol=2e-6, atol=1e-7)


def test_earlyfusion_class_matches_reference_golden(golden, workdir):
    import acoss_b200
    g = golden("earlyfusion_reference.npz")
    csv, fdir, ds = _early_dataset(workdir)
    e = acoss_b200.EarlyFusion(csv, fdir, shortname="g")
    for i, f in enumerate(ds["block_feats"]):
        e.set_block_features(i, f)
    e.all_pairwise(1, symmetric=True)
    for s in ("mfccs", "ssms", "chromas", "early"):
        assert np.array_equal(np.array(e.Ds[s]), g["D_" + s]), s
    e.do_late_fusion()
    for s in ("late", "early+late"):
        np.testing.assert_allclose(np.array(e.Ds[s]), g["D_" + s], rtol=2e-5, atol=1e-9)
    for s in ("mfccs", "ssms", "chromas", "early", "late", "early+late"):
        assert list(e.getEvalStatistics(s)[:4]) == g["eval_" + s].tolist()
    e.cleanup_memmap()
    # the reference's cache files now exist: the benchmark entry point runs from them (both names)
    res = acoss_b200.benchmark(csv, fdir, algorithm="EarlyFusionTraile", shortname="g")
    assert list(res["early+late"][:4]) == g["eval_early+late"].tolist()


def test_earlyfusion_errors(ctx):
    import acoss_b200
    from acoss_b200 import synth
    ds = synth.make_block_features(1, 2, 0, 3, 12, 14)
    ctx.earlyfusion_upload(ds["block_feats"])
    with pytest.raises(acoss_b200.AcossError):
        ctx.earlyfusion_pairs(np.array([[0, 5]], np.int32))
    with pytest.raises(acoss_b200.AcossError):
        ctx.earlyfusion_pairs(np.array([[0, 1]], np.int32), kappa=100)       # kth out of bounds
    sc = ctx.earlyfusion_pairs(np.array([[0, 1]], np.int32), kappa=0)         # kappa = 0: all ones
    M, N = ds["block_feats"][0]["mfccs"].shape[0], ds["block_feats"][1]["mfccs"].shape[0]
    assert sc[0].tolist() == [float(min(M, N) - 3)] * 4


def test_free_functions_vs_ports(ctx, oracle):
    """The reference's free functions (acoss/algorithms/utils/{alignment_tools,cross_recurrence,similarity_fusion}.py) under
    acoss_b200.utils, each through the C-ABI, against the numpy restatements of oracle/ports.py (pinned against the reference's own
    outputs in tests/test_oracle.py).  float32 GEMM-shaped results: rtol 2e-5 (summation order); selections: exact."""
    from acoss_b200 import utils as U
    from oracle import ports as P
    U.set_default_context(ctx)
    try:
        rng = np.random.default_rng(11)
        X = rng.random((70, 36)).astype(np.float32)
        Y = rng.random((93, 36)).astype(np.float32)
        np.testing.assert_allclose(U.get_csm(X, Y), P.get_csm(X, Y), rtol=2e-5, atol=2e-6)
        np.testing.assert_allclose(U.get_csm_cosine(X, Y), P.get_csm_cosine(X, Y), rtol=2e-5, atol=2e-6)
        S = U.get_ssm(X)
        assert S.shape == (70, 70) and np.all(np.diag(S) == 0)
        np.testing.assert_allclose(S[np.triu_indices(70, 1)], P.get_csm(X, X)[np.triu_indices(70, 1)], rtol=2e-5, atol=2e-5)
        C1 = rng.random(12).astype(np.float32)
        for sh in (0, 3, 11):
            C2 = np.roll(C1, sh) + 0.01 * rng.random(12).astype(np.float32)
            assert U.get_oti(C1, C2) == P.get_oti(C1, C2) == sh
        C2 = np.roll(C1, 5)
        for fn_u, fn_p in ((U.get_csm_cosine, P.get_csm_cosine), (U.get_csm, P.get_csm)):
            got = U.get_csm_blocked_oti(X, Y, C1, C2, fn_u)
            oti = P.get_oti(C1, C2)
            X1 = np.roll(X.reshape(70, 3, 12), oti, axis=2).reshape(70, 36)
            np.testing.assert_allclose(got, fn_p(X1, Y), rtol=2e-5, atol=2e-6)
        # selections on a matrix without ties
        D = P.get_csm(X, Y)
        for kappa in (0, 0.1, 0.25, 7):
            got = U.csm_to_binary(D, kappa)
            exp = P.csm_to_binary(D, kappa)
            assert got.dtype == np.uint8 and np.array_equal(got, np.asarray(exp, dtype=np.uint8)), kappa
        import acoss_b200
        with pytest.raises(acoss_b200.AcossError):
            U.csm_to_binary(D, 93)
        np.testing.assert_allclose(U.getWCSM(D, 5, 8), P.getWCSM(D, 5, 8), rtol=2e-5, atol=1e-7)
        B = (rng.random((40, 57)) < 0.3).astype(np.uint8)
        assert float(np.float32(U.smith_waterman_constrained(B))) == float(np.float32(oracle.sw_constrained(B)))
        # similarity fusion stages, float64 (exact arithmetic order: 1e-12) and float32
        n = 48
        Ds = []
        for _ in range(3):
            A = rng.random((n, 6))
            Ds.append(np.sqrt(((A[:, None, :] - A[None, :, :]) ** 2).sum(2)))
        for dt, tol in ((np.float64, 1e-12), (np.float32, 2e-5)):
            Dk = [d.astype(dt) for d in Ds]
            Ws = [P.getW(np.array(d), 5) for d in Dk]
            for d, w in zip(Dk, Ws):
                got = U.getW(d, 5)
                assert got.dtype == dt
                np.testing.assert_allclose(got, w, rtol=tol, atol=1e-12)
            np.testing.assert_allclose(U.getP(Ws[0]), P.getP(Ws[0]), rtol=tol)
            Sg = U.getS(Ws[1], 5)
            assert Sg.shape == (n, n) and np.all(np.diff(Sg.indptr) == 5)
            np.testing.assert_allclose(Sg.toarray(), P.getS(Ws[1], 5), rtol=tol)
            np.testing.assert_allclose(U.doSimilarityFusionWs(Ws, K=5, niters=4, reg_diag=1), P.snf_ws(Ws, 5, 4, 1), rtol=10 * tol)
        Wsg, fused = U.doSimilarityFusion(Ds, K=5, niters=3)
        Wsp, fusedp = P.snf(Ds, 5, 3, 1)
        np.testing.assert_allclose(fused, fusedp, rtol=1e-11)
    finally:
        U.set_default_context(None)


def test_serra09_topk_vs_full_matrix(ctx):
    """acoss_serra09_topk (query x database, per-query top-k kept on the device) against a stable argsort of the full rectangle
    computed pair by pair through acoss_serra09_pairs and normalised the way Serra09 / ChenFusion.normalize_by_length do
    (float64 arithmetic, float32 store): indices and scores exact.  The database repeats tracks (exact ties -> lowest position),
    spans two device chunks (5000 > 4096 - k) and contains the queries themselves (left out of their own ranking)."""
    from acoss_b200 import synth
    import acoss_b200
    labels, works, tids, hp = synth.make_tracks(10, 4, 0, 7, 40, 90)
    ctx.upload_tracks(hp)
    n = len(hp)
    lens = np.array([len(h) for h in hp], np.float64)
    rng = np.random.default_rng(3)
    queries = np.array([0, 17, 39], np.int32)
    db = rng.integers(0, n, 5000).astype(np.int32)
    prm = ctx.params()
    pairs = np.stack([np.repeat(queries, len(db)), np.tile(db, len(queries))], 1).astype(np.int32)
    q, d = ctx.serra09_pairs(pairs, prm, want_dmax=True)
    for which, norm, raw in (("qmax", "serra09", q), ("dmax", "chen", d), ("qmax", None, q)):
        full = raw.reshape(len(queries), len(db)).astype(np.float32)
        if norm == "serra09":
            full = (full.astype(np.float64) / np.sqrt(lens[db])[None, :]).astype(np.float32)
        elif norm == "chen":
            with np.errstate(divide="ignore"):
                full = (np.sqrt(lens[db])[None, :] / full.astype(np.float64)).astype(np.float32)
        for k in (1, 10, 1000):
            idx, sc = ctx.serra09_topk(queries, db, k, prm, which=which, norm=norm)
            assert idx.shape == (3, k) and sc.shape == (3, k)
            for a, qt in enumerate(queries):
                keep = np.flatnonzero(db != qt)
                row = full[a, keep]
                order = np.argsort(row if norm == "chen" else -row, kind="stable")[:k]
                assert np.array_equal(idx[a, :len(order)], keep[order]), (which, norm, k, a)
                assert np.array_equal(sc[a, :len(order)], row[order])
    # fewer database entries than k: -1 / -inf fill; argument errors
    idx, sc = ctx.serra09_topk([1], [1, 2, 3], 5, prm)
    assert list(idx[0, 2:]) == [-1, -1, -1] and np.all(np.isneginf(sc[0, 2:])) and set(idx[0, :2]) == {1, 2}
    with pytest.raises(acoss_b200.AcossError):
        ctx.serra09_topk([0], [1], 1001, prm)
    with pytest.raises(acoss_b200.AcossError):
        ctx.serra09_topk([n], [1], 3, prm)


def test_caller_stream(ctx, oracle):
    """acoss_ctx_set_stream: the context works on a stream the caller owns (a torch stream here): device-output scores written by
    acoss_serra09_pairs_dev are visible to a torch kernel enqueued on the same stream without any synchronisation in between."""
    import torch
    rng = np.random.default_rng(8)
    tracks = _tracks(rng, [70, 95, 120, 64])
    ctx.upload_tracks(tracks)
    pairs = np.array([(0, 1), (0, 2), (1, 3), (2, 3), (1, 2)], np.int32)
    exp, _, _ = oracle.serra09_pairs(tracks, pairs)
    st = torch.cuda.Stream(device=0)
    out = torch.zeros(len(pairs), dtype=torch.float32, device="cuda:0")
    try:
        ctx.set_stream(st.cuda_stream)
        with torch.cuda.stream(st):
            ctx.serra09_pairs_dev(pairs, ctx.params(), out.data_ptr(), 0)
            doubled = out * 2.0                                  # same stream: ordered after the library's kernels
        st.synchronize()
        assert np.array_equal(out.cpu().numpy(), exp)
        assert np.array_equal(doubled.cpu().numpy(), 2.0 * exp)
    finally:
        ctx.set_stream(None)
    q, _ = ctx.serra09_pairs(pairs)                              # back on the context's own stream
    assert np.array_equal(q, exp)
