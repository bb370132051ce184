"""CPU tests of the oracle: KATs, the two independent restatements
against each other, the golden fixtures frozen from the reference, and -- when
/root/reference is present -- the reference's own code run verbatim."""
import numpy as np
import pytest


def test_kat_identity_gap_allones(oracle):
    L = 12
    I = np.eye(L, dtype=np.uint8)
    assert oracle.cover_song_similarity(I)[1] == L - 2
    assert oracle.sw_constrained(I) == L - 3
    G = I.copy()
    G[6, 6] = 0
    assert oracle.cover_song_similarity(G)[1] == 8.5
    assert oracle.cover_song_similarity(G, "chen17")[1] == 8.5
    Z = np.zeros((L, L), np.uint8)
    assert oracle.cover_song_similarity(Z)[1] == 0 and oracle.cover_song_similarity(Z, "chen17")[1] == 0
    assert oracle.sw_constrained(Z) == 0
    O1 = np.ones((L, L), np.uint8)
    assert oracle.cover_song_similarity(O1)[1] == 10.0
    assert oracle.cover_song_similarity(O1, "chen17")[1] == 14.0
    assert oracle.sw_constrained(O1) == 9.0
    assert oracle.sw_constrained(np.ones((3, 20), np.uint8)) == 0.0


def test_kat_oti_roll(oracle):
    rng = np.random.default_rng(3)
    q = rng.random((60, 12)).astype(np.float32)
    for s in range(12):
        r = np.roll(q, s, axis=1)
        o = oracle.oti(q, r)
        assert o % 12 == (12 - s) % 12
        R1 = oracle.chroma_cross_similarity(q, r)
        R0 = oracle.chroma_cross_similarity(q, q)
        assert (R1 == R0).all()


def test_c_vs_numpy_restatement(oracle):
    from oracle import numpy_ref as NR
    rng = np.random.default_rng(0)
    for (nq, nr, m, tau) in [(40, 37, 9, 1), (25, 31, 5, 2), (14, 12, 1, 1), (30, 30, 3, 3)]:
        q = rng.random((nq, 12)).astype(np.float32)
        r = (np.roll(q, 5, axis=1)[: nr] if nr <= nq else rng.random((nr, 12))).astype(np.float32)
        r = (r + 0.1 * rng.random(r.shape)).astype(np.float32)
        R, info = oracle.chroma_cross_similarity(q, r, m=m, tau=tau, full=True)
        R2, D2, tq2, tr2 = NR.crp(q, r, m=m, tau=tau)
        assert info["oti"] == NR.oti(q, r)
        assert np.array_equal(info["dist"], D2)
        assert np.array_equal(info["thr_q"], tq2) and np.array_equal(info["thr_r"], tr2)
        assert np.array_equal(R, R2)
        for a in ("serra09", "chen17"):
            for (go, ge) in [(0.5, 0.5), (0.5, 0.7), (0.3, 0.9)]:
                assert oracle.cover_song_similarity(R, a, go, ge)[1] == NR.cover_similarity(R, a, go, ge)[1]


def test_percentile_equals_numpy_linear(oracle, golden):
    g = golden("percentile_kat.npz")
    for k in range(int(g["n"])):
        v = g["v%d" % k]
        p = oracle.percentile(v, 0.095)
        assert p == g["p"][k]
        ref = np.percentile(v.astype(np.float64), 9.5)
        assert abs(p - ref) <= 2e-6 * max(1.0, abs(ref))


def test_too_short_raises(oracle):
    q = np.ones((9, 12), np.float32)
    with pytest.raises(RuntimeError):
        oracle.chroma_cross_similarity(q, q)
    assert oracle.embedded_len(10) == 1 and oracle.embedded_len(9) == 0


def test_sw_oracle_vs_reference_golden(oracle, golden):
    g = golden("sw_reference.npz")
    for k in range(int(g["n"])):
        B = g["B%d" % k]
        assert np.float32(oracle.sw_constrained(B)) == np.float32(g["scores"][k])


def test_eval_restatement_vs_reference_golden(oracle, golden):
    g = golden("eval_reference.npz")
    for c in range(int(g["n"])):
        D, labels, res = g["D%d" % c], [str(x) for x in g["labels%d" % c]], g["res%d" % c]
        cl = {}
        for i, l in enumerate(labels):
            cl.setdefault(l, set([])).add(i)
        MR, MRR, MDR, MAP, tops = oracle.eval_statistics(D, cl)
        if not int(g["ties%d" % c]):
            assert [MR, MRR, MDR, MAP] == list(res[:4])
            assert list(tops) == list(res[4:])
        else:   # ties: the reference's unstable argsort may order equal scores differently
            assert abs(MAP - res[3]) < 0.1


def test_serra09_golden_consistent_with_oracle(oracle, golden):
    g = golden("serra09_small.npz")
    pairs = g["crp_pairs"]
    for k, (a, b) in enumerate(pairs):
        R, info = oracle.chroma_cross_similarity(g["ds%d" % a], g["ds%d" % b], full=True)
        assert np.array_equal(R, g["crp%d" % k])
        assert info["oti"] == int(g["oti%d" % k])
        q = oracle.cover_song_similarity(R)[1]
        i, j = (a, b) if a < b else (b, a)
        if a < b:
            assert q == g["serra_raw"][a, b]
    # downsampling restatement == what the reference class cached
    for i in range(int(g["n_tracks"])):
        assert np.array_equal(oracle.sync_median(g["hpcp%d" % i], int(g["fac"])), g["ds%d" % i])
    lengths = g["lengths"]
    assert np.array_equal(oracle.serra09_normalize(g["serra_raw"], lengths), g["serra_norm"])
    with np.errstate(divide="ignore"):
        assert np.array_equal(oracle.chen_normalize(g["chen_q_raw"], lengths), g["chen_q_norm"])


def test_reference_verbatim_when_present(oracle):
    from oracle import refload
    if not refload.available():
        pytest.skip("/root/reference not present (GPU box)")
    ns = refload.load()
    rng = np.random.default_rng(11)
    for _ in range(5):
        B = (rng.random((int(rng.integers(4, 90)), int(rng.integers(4, 90)))) < 0.2).astype(np.uint8)
        assert np.float32(ns.align.smith_waterman_constrained(B)) == np.float32(oracle.sw_constrained(B))


def test_ports_vs_reference_goldens(oracle, golden):
    """oracle/ports.py (the numpy ports that bench.py times as the CPU baseline of C3 / C4 / C5 on the GPU box, where the
    reference itself is absent) against the goldens made from the reference run verbatim."""
    from oracle import ports as P
    from oracle.make_golden import early_dataset
    # EarlyFusion: matrices and scores of the small golden pairs
    g = golden("earlyfusion_reference.npz")
    bf = early_dataset()["block_feats"]
    for k, (a, b) in enumerate(g["mat_pairs"]):
        sc, C, B = P.early_pair(bf[a], bf[b], 0.1, 10, full=True)
        np.testing.assert_allclose(np.stack(C), g["csm%d" % k], rtol=1e-5, atol=1e-6)
        assert np.array_equal(np.stack(B), g["bin%d" % k])
        assert sc.astype(np.float32).tolist() == [g["D_" + s][a, b] for s in ("mfccs", "ssms", "chromas", "early")]
        assert P.get_oti(bf[a]["chroma_med"], bf[b]["chroma_med"]) == int(g["oti%d" % k])
    # SiMPle at both golden sizes
    g = golden("simple_ftm2d_reference.npz")
    feats = [g["simple_feat%d" % i] for i in range(10)]
    got = [P.simple_pair(feats[a], feats[b]) for a, b in g["simple_pairs"]]
    np.testing.assert_allclose(got, g["simple_scores"], rtol=1e-9, atol=1e-11)
    g2 = golden("simple_large.npz")
    f2 = [g2["feat%d" % i] for i in range(int(g2["n"]))]
    got = [P.simple_pair(f2[a], f2[b]) for a, b in g2["pairs"][:12]]
    np.testing.assert_allclose(got, g2["scores"][:12], rtol=1e-9, atol=1e-11)
    # FTM2D
    S = [g["ftm_shingle%d" % i] for i in range(10)]
    assert np.float32(P.ftm2d_pair(S[0], S[3])) == g["ftm_D"][0, 3]
    # SNF incl. the reference's list aliasing
    gs = golden("snf_reference.npz")
    for c in range(int(gs["n"])):
        nm, N, K, nit = (int(x) for x in gs["cfg%d" % c])
        if N > 128:
            continue
        mats = [gs["D%d_%d" % (c, k)] for k in range(nm)]
        np.testing.assert_allclose(P.snf(mats, K=K, niters=nit, reg_diag=1)[1], gs["fused%d" % c], rtol=1e-9, atol=1e-12)
