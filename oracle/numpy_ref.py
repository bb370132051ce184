# -*- coding: utf-8 -*-
"""
CPU ORACLE (second, independent restatement) -- TEST INFRASTRUCTURE ONLY.

A numpy / pure-Python restatement of the same SURVEY.md Appendix A decisions as
``acoss_oracle.c``, written independently (different data layout, numpy's own
``np.percentile``/``np.roll``), used only to cross-check the C oracle on small
cases.  PARITY UNPINNED for the same reason as the C file (essentia absent).

Reference call sites restated: acoss/algorithms/rqa_serra09.py:60-67,
acoss/algorithms/latefusion_chen.py:63-71.
"""
import numpy as np

f32 = np.float32


def _seq_dot(a, b):
    """sequential fp32 sum of fp32 products (D5: no FMA, left to right)."""
    acc = f32(0.0)
    for x, y in zip(a, b):
        acc = f32(acc + f32(x * y))
    return acc


def global_chroma(x):
    g = np.zeros(12, f32)
    for fr in np.asarray(x, f32):
        g = (g + fr).astype(f32)
    mx = g.max()
    if mx != 0:
        g = (g / mx).astype(f32)
    return g


def oti(q, r):
    """D1: first argmax over k=0..12 of <g_q, np.roll(g_r, k)>."""
    gq, gr = global_chroma(q), global_chroma(r)
    vals = [_seq_dot(gq, np.roll(gr, k)) for k in range(13)]
    return int(np.argmax(np.array(vals, f32)))


def stack(x, m, tau):
    """D3/D4: rows i = 0, tau, ... < n - m*tau; m == 1 -> unchanged."""
    x = np.asarray(x, f32)
    if m == 1:
        return x
    n = x.shape[0]
    inc = m * tau
    if n < inc + 1:
        raise RuntimeError("too short")
    rows = []
    for i in range(0, n - inc, tau):
        rows.append(np.concatenate([x[s] for s in range(i, i + inc, tau)]))
    return np.array(rows, f32)


def percentile(v, kappa):
    """D6.  fp32 rank arithmetic; equals np.percentile(method='linear') up to fp32."""
    s = np.sort(np.asarray(v, f32))
    n = len(s)
    if n == 1:
        return s[0]
    q = f32(f32(f32(kappa) * f32(100.0)) / f32(100.0))
    k = f32(f32(n - 1) * q)
    fl, ce = np.floor(k), np.ceil(k)
    if fl == ce:
        return s[int(fl)]
    d0 = f32(s[int(fl)] * f32(ce - k))
    d1 = f32(s[int(ce)] * f32(k - fl))
    return f32(d0 + d1)


def crp(q, r, m=9, tau=1, kappa=0.095, use_oti=True):
    q = np.asarray(q, f32)
    r = np.asarray(r, f32)
    if use_oti:
        o = oti(q, r)
        r = np.roll(r, o, axis=1)
    X, Y = stack(q, m, tau), stack(r, m, tau)
    M, N = len(X), len(Y)
    xx = [_seq_dot(x, x) for x in X]
    yy = [_seq_dot(y, y) for y in Y]
    D = np.zeros((M, N), f32)
    for i in range(M):
        for j in range(N):
            item = f32(f32(xx[i] - f32(f32(2.0) * _seq_dot(X[i], Y[j]))) + yy[j])
            D[i, j] = np.sqrt(max(item, f32(0.0)), dtype=f32)
    tq = np.array([percentile(D[i, :], kappa) for i in range(M)], f32)
    tr = np.array([percentile(D[:, j], kappa) for j in range(N)], f32)
    R = ((D <= tq[:, None]) & (D <= tr[None, :])).astype(np.uint8)
    return R, D, tq, tr


def cover_similarity(R, alignment="serra09", go=0.5, ge=0.5):
    """D8/D9 literal transcription; returns (score matrix f32, max)."""
    R = np.asarray(R)
    M, N = R.shape
    S = np.zeros((M, N), f32)
    g = lambda v: f32(go) if v == 1 else f32(ge)
    for i in range(2, M):
        for j in range(2, N):
            p1 = S[i - 1, j - 1]
            p2 = S[i - 2, j - 1]
            p3 = S[i - 1, j - 2]
            if alignment == "chen17":
                p2 = f32(p2 + f32(R[i - 1, j]))
                p3 = f32(p3 + f32(R[i, j - 1]))
            if R[i, j] == 1:
                S[i, j] = f32(max(p1, p2, p3) + f32(1.0))
            else:
                S[i, j] = max(f32(0.0), f32(p1 - g(R[i - 1, j - 1])), f32(p2 - g(R[i - 2, j - 1])),
                              f32(p3 - g(R[i - 1, j - 2])))
    return S, float(S.max()) if S.size else 0.0
