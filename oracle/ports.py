# -*- coding: utf-8 -*-
"""
CPU ORACLE (ports) -- TEST INFRASTRUCTURE ONLY.

numpy restatements of the reference's in-tree per-pair functions for EarlyFusionTralie, SiMPle, FTM2D and the
N x N similarity network fusion.  /root/reference does not exist on the GPU box, so the CPU baselines of
`bench.py --config C3|C4|C5` (cpu_baseline.kind = "port") and the `--impl reference` arm time THESE; they are pinned
against the goldens generated from the reference run verbatim (tests/test_oracle.py::test_ports_vs_reference_goldens).
Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this module.

Each function cites the reference lines it follows (paths relative to the reference root).
"""
import numpy as np

from . import oracle as O


# ---- acoss/algorithms/utils/cross_recurrence.py -------------------------------------------------------------------
def get_csm(X, Y):
    """Euclidean cross-similarity by Gram expansion, negatives clipped (cross_recurrence.py:30-48).  float32 in -> float32."""
    C = np.sum(X ** 2, 1)[:, None] + np.sum(Y ** 2, 1)[None, :] - 2 * X.dot(Y.T)
    C[C < 0] = 0
    return np.sqrt(C)


def get_csm_cosine(X, Y):
    """1 - cosine similarity, zero norms replaced by 1 (cross_recurrence.py:53-73)."""
    xn = np.sqrt(np.sum(X ** 2, 1))
    xn[xn == 0] = 1
    yn = np.sqrt(np.sum(Y ** 2, 1))
    yn[yn == 0] = 1
    return 1 - (X / xn[:, None]).dot((Y / yn[:, None]).T)


def get_oti(C1, C2):
    """argmax_k sum(roll(C1, k) * C2), first maximum (cross_recurrence.py:75-103)."""
    n = len(C1)
    s = np.zeros(n)
    for k in range(n):
        s[k] = np.sum(np.roll(C1, k) * C2)
    return int(np.argmax(s))


def get_csm_blocked_oti(X, Y, C1, C2, csm_fn=get_csm_cosine):
    """roll the 12 bins of every chroma frame of every block of X by the global OTI, then the CSM
    (cross_recurrence.py:105-134)."""
    nb = len(C1)
    per = X.shape[1] // nb
    X1 = np.roll(X.reshape(X.shape[0], per, nb), get_oti(C1, C2), axis=2).reshape(X.shape[0], per * nb)
    return csm_fn(X1, Y)


def csm_to_binary(D, kappa):
    """row-wise: the int(round(kappa * n_cols)) smallest entries of every row -> 1 (cross_recurrence.py:136-161;
    np.round is half-to-even)."""
    M, N = D.shape
    if kappa == 0:
        return np.ones((M, N), np.uint8)
    k = int(np.round(kappa * N)) if kappa < 1 else int(kappa)
    J = np.argpartition(D, k, 1)[:, :k]
    B = np.zeros((M, N), np.uint8)
    B[np.arange(M)[:, None], J] = 1
    return B


# ---- acoss/algorithms/utils/similarity_fusion.py ------------------------------------------------------------------
def getWCSM(C, k1, k2, Mu=0.5):
    """similarity_fusion.py:38-54: row mean of the k2 smallest, column mean of the k1 smallest, eps = (r + c + C) / 3,
    W = exp(-C^2 / (2 (Mu eps)^2))."""
    m1 = np.mean(np.partition(C, k2, 1)[:, :k2], 1)
    m2 = np.mean(np.partition(C, k1, 0)[:k1, :], 0)
    eps = (m1[:, None] + m2[None, :] + C) / 3
    return np.exp(-C ** 2 / (2 * (Mu * eps) ** 2))


def getW(D, K, Mu=0.5):
    """similarity_fusion.py:15-36."""
    S = 0.5 * (D + D.T)
    np.fill_diagonal(S, 0)
    md = np.mean(np.partition(S, K + 1, 1)[:, :K + 1], 1) * float(K + 1) / float(K)
    eps = (md[:, None] + md[None, :] + S) / 3
    den = 2 * (Mu * eps) ** 2
    den[den == 0] = 1
    return np.exp(-S ** 2 / den)


def getP(W):
    """similarity_fusion.py:101-122 (diagRegularize=False branch, the only one used)."""
    rs = np.sum(W, 1)
    rs[rs == 0] = 1
    return W / rs[:, None]


def getS(W, K):
    """similarity_fusion.py:124-144, as a dense (N, N) matrix with K non-zeros per row."""
    N = W.shape[0]
    J = np.argpartition(-W, K, 1)[:, :K]
    V = W[np.arange(N)[:, None], J]
    sn = np.sum(V, 1)
    sn[sn == 0] = 1
    S = np.zeros((N, N), W.dtype if W.dtype == np.float64 else np.float64)
    S[np.arange(N)[:, None], J] = V / sn[:, None]
    return S


def snf_ws(Ws, K=5, niters=20, reg_diag=1):
    """doSimilarityFusionWs (similarity_fusion.py:146-186) INCLUDING its list aliasing: after the first
    iteration `Pts` and `nextPts` are one list, so matrix i sees the already-updated matrices k < i."""
    Ps = [getP(W) for W in Ws]
    Ss = [getS(W, K) for W in Ws]
    Pts = [np.array(P, dtype=np.float64) for P in Ps]
    nxt = [np.zeros(P.shape) for P in Pts]
    n = len(Pts)
    pix = np.arange(Ws[0].shape[0])
    for _ in range(niters):
        for i in range(n):
            nxt[i] *= 0
            for k in range(n):
                if k != i:
                    nxt[i] += Pts[k]
            nxt[i] /= float(n - 1)
            nxt[i] = Ss[i].dot((Ss[i].dot(nxt[i].T)).T)
            if reg_diag > 0:
                nxt[i][pix, pix] += reg_diag
        Pts = nxt
    out = np.zeros(Pts[0].shape)
    for P in Pts:
        out += P
    return out / n


def snf(mats, K=5, niters=5, reg_diag=1):
    """doSimilarityFusion (similarity_fusion.py:188-196) -> (Ws, fused)."""
    Ws = [getW(np.array(D), K) for D in mats]
    return Ws, snf_ws(Ws, K, niters, reg_diag)


# ---- acoss/algorithms/earlyfusion_traile.py:157-198 ----------------------------------------------------------------
def early_pair(f1, f2, kappa=0.1, K=10, full=False):
    """EarlyFusion.similarity for one pair of block-feature dicts: 4 scores (mfccs, ssms, chromas, early); the constrained
    Smith-Waterman is the C oracle's (pinned against the reference's numba code)."""
    C = [get_csm(f1["mfccs"], f2["mfccs"]), get_csm(f1["ssms"], f2["ssms"]),
         get_csm_blocked_oti(f1["chromas"], f2["chromas"], f1["chroma_med"], f2["chroma_med"])]
    Wsum = np.zeros_like(C[0])
    for c in C:
        Wsum += getWCSM(c, K, K)
    C.append(np.exp(-Wsum))
    B = [csm_to_binary(c, kappa) for c in C]
    scores = np.array([O.sw_constrained(b) for b in B], np.float64)
    return (scores, C, B) if full else scores


# ---- acoss/algorithms/simple_silva.py:45-54, 68-118 ----------------------------------------------------------------
def simple_pair(Si, Sj, sslen=10):
    """-median_i min_j sum_{c, k < sslen} (a[c, i + k] - b[c, j + k])^2 with b = Sj rolled by the OTI of the time-summed
    profiles (argsort, last index).  Evaluated directly (Gram + sliding diagonal sums) in float64; the reference's
    FFT-seeded update gives the same values to ~1e-13."""
    pa, pb = np.sum(Si, 1), np.sum(Sj, 1)
    v = np.array([np.dot(pa, np.roll(pb, k)) for k in range(12)])
    b = np.roll(Sj, int(np.argsort(v)[-1]), axis=0)
    a = Si
    La, Lb = a.shape[1], b.shape[1]
    na, nb = La - sslen + 1, Lb - sslen + 1
    G = a.T.dot(b)                                              # (La, Lb) frame Gram
    cs = np.zeros((La + 1, Lb + 1))
    # diagonal prefix sums: cs[i + 1, j + 1] = cs[i, j] + G[i, j]
    for i in range(La):
        cs[i + 1, 1:] = cs[i, :-1] + G[i]
    dot = cs[sslen:sslen + na, sslen:sslen + nb] - cs[:na, :nb]
    a2 = np.concatenate([[0.0], np.cumsum(np.sum(a * a, 0))])
    b2 = np.concatenate([[0.0], np.cumsum(np.sum(b * b, 0))])
    d = (a2[sslen:] - a2[:na])[:, None] + (b2[sslen:] - b2[:nb])[None, :] - 2 * dot
    return -float(np.median(np.min(d, 1)))


# ---- acoss/algorithms/ftm2d.py:85-97 --------------------------------------------------------------------------------
def ftm2d_pair(s1, s2):
    return float(np.exp(-np.sum((s1 - s2) ** 2)))


# ---- per-track features (acoss/algorithms/ftm2d.py:52-66, 100-138; simple_silva.py:34-43, 56-66) ---------------------
def sync_median(data, idx):
    """librosa.util.sync(data, idx, aggregate=np.median) (librosa 0.6.1): segments between the unique boundaries
    {0} u idx u {T} along the last axis."""
    T = data.shape[-1]
    b = np.unique(np.concatenate([[0], np.asarray(idx, dtype=int), [T]])).astype(int)
    b = b[(b >= 0) & (b <= T)]
    out = np.empty(data.shape[:-1] + (len(b) - 1,), dtype=data.dtype)
    for k in range(len(b) - 1):
        out[..., k] = np.median(data[..., b[k]:b[k + 1]], axis=-1)
    return out


def chrompwr(X, P=.5):
    """ftm2d.py:100-117: chroma columns to the power P, norms preserved."""
    n = np.sqrt(np.sum(X * X, axis=0))[None, :].repeat(X.shape[0], 0)
    n[n == 0] = 1
    Xp = np.power(X / n, P)
    npw = np.sqrt(np.sum(Xp * Xp, axis=0))[None, :].repeat(X.shape[0], 0)
    npw[npw == 0] = 1.
    return n * (Xp / npw)


def ftm2d_shingle(chroma_T, onsets, PWR=1.96, WIN=75, C=5):
    """FTM2D.load_features (ftm2d.py:52-66) for one track; chroma_T: (12, T) float32.  scipy.fftpack.fft2 keeps the float32
    patch in single precision (complex64), like the reference."""
    import scipy.fftpack
    hp = sync_median(chroma_T, onsets)
    ch = chrompwr(hp, PWR)
    nb = ch.shape[1]
    if nb < WIN:
        raise ValueError("fewer beats than WIN")
    mat = np.zeros((12 * WIN, nb - WIN + 1))
    for i in range(nb - WIN + 1):
        F = scipy.fftpack.fft2(ch[:, i:i + WIN])
        mat[:, i] = scipy.fftpack.fftshift(np.sqrt(np.real(F) ** 2 + np.imag(F) ** 2)).flatten()
    sh = mat.T
    nrm = np.sqrt(np.sum(sh ** 2, 1))
    nrm[nrm == 0] = 1
    sh = np.log(C * sh / nrm[:, None] + 1)
    s = np.median(sh, 0)
    return s / np.sqrt(np.sum(s ** 2))


def simple_features(chroma_T, WIN=200, SKIP=100, win_len_smooth=4):
    """Simple.load_features + Simple.smooth (simple_silva.py:34-43, 56-66) for one track; chroma_T: (12, T)."""
    nf = np.zeros((chroma_T.shape[0], int(chroma_T.shape[1] / SKIP)))
    for k in range(nf.shape[1]):
        nf[:, k] = np.mean(chroma_T[:, k * SKIP:k * SKIP + WIN], axis=1)
    if nf.shape[1] == 0:
        return nf                                                       # a track shorter than SKIP: no columns (np.convolve refuses empty input)
    n = win_len_smooth + 2
    win = 0.5 - 0.5 * np.cos(2.0 * np.pi * np.arange(n) / (n - 1))       # symmetric Hann (fftbins=False)
    win = win / np.sum(win)
    out = np.stack([np.convolve(r, win, mode='full') for r in nf])
    st = (len(win) - 1) // 2
    out = out[:, st:st + nf.shape[1]]
    length = np.sqrt(np.sum(np.abs(out) ** 2, axis=0, keepdims=True))
    length[length < np.finfo(out.dtype).tiny] = 1.0
    return out / length


# ---- acoss/algorithms/earlyfusion_traile.py:104-150, 214-247 ---------------------------------------------------------
def resize_block(X, i1, i2, frames_per_block):
    """resize_block (earlyfusion_traile.py:214-247, median_aggregate=False) = skimage.transform.resize(X[i1:i2].astype(float64),
    (frames_per_block, n_feats), anti_aliasing=True, mode='constant').  skimage is absent from this image (UNPINNED); restated
    from skimage 0.15 / 0.16: (1) scipy.ndimage.gaussian_filter with sigma = max(0, (factor - 1) / 2) per axis, mode='constant',
    cval=0 -- the REAL scipy routine is used here; (2) order-1 warp sampling source row factor * (o + 0.5) - 0.5, neighbours
    outside the image = 0; the final clip to the input range is a no-op for bilinear weights."""
    import scipy.ndimage as ndi
    x = np.asarray(X[i1:i2, :], dtype=np.float64)
    n_in, n_out = x.shape[0], int(frames_per_block)
    f = float(n_in) / float(n_out)
    sig = max(0.0, (f - 1.0) / 2.0)
    g = ndi.gaussian_filter(x, (sig, 0.0), mode='constant', cval=0.0)
    r = f * (np.arange(n_out) + 0.5) - 0.5
    rf, rc = np.floor(r).astype(int), np.ceil(r).astype(int)
    dr = r - rf

    def rows(idx):
        ok = (idx >= 0) & (idx <= n_in - 1)
        out = np.zeros((n_out, x.shape[1]))
        out[ok] = g[idx[ok]]
        return out
    ret = (1.0 - dr)[:, None] * rows(rf) + np.where(rc != rf, dr, 0.0)[:, None] * rows(rc)
    ret[np.isinf(ret)] = 0
    ret[np.isnan(ret)] = 0
    return ret


def early_block_features(mfcc_htk, chroma, onsets, blocksize=20, mfccs_per_block=50, chromas_per_block=40):
    """EarlyFusion.load_features (earlyfusion_traile.py:104-150) for one track: mfcc_htk (n_mfcc, T), chroma (T, 12), onsets."""
    mfcc = np.array(mfcc_htk).T
    mfcc[np.isnan(mfcc)] = 0
    n_blocks = len(onsets) - blocksize
    bf = {'mfccs': np.zeros((n_blocks, mfccs_per_block * mfcc.shape[1]), np.float32)}
    pix = np.arange(mfccs_per_block)
    I, J = np.meshgrid(pix, pix)
    bf['ssms'] = np.zeros((n_blocks, mfccs_per_block * (mfccs_per_block - 1) // 2), np.float32)
    for b in range(n_blocks):
        x = resize_block(mfcc, onsets[b], onsets[b + blocksize - 1], mfccs_per_block)
        x -= np.mean(x, 0)[None, :]
        xnorm = np.sqrt(np.sum(x ** 2, 1))[:, None]
        xnorm[xnorm == 0] = 1
        xn = x / xnorm
        bf['mfccs'][b, :] = xn.flatten()
        sq = np.sum(xn ** 2, 1)                                  # get_ssm (cross_recurrence.py:10-28)
        D = sq[:, None] + sq[None, :] - 2 * xn.dot(xn.T)
        D[D < 0] = 0
        np.fill_diagonal(D, 0)
        bf['ssms'][b, :] = np.sqrt(D)[I < J]
    bf['chromas'] = np.zeros((n_blocks, chromas_per_block * chroma.shape[1]), np.float32)
    bf['chroma_med'] = np.median(chroma, axis=0)
    for b in range(n_blocks):
        bf['chromas'][b, :] = resize_block(chroma, onsets[b], onsets[b + blocksize], chromas_per_block).flatten()
    return bf
