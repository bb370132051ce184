#!/usr/bin/env python
"""Throughput of the OTHER BASELINE.json configs (C3 EarlyFusionTralie, C4 LateFusionChen, C5 SiMPle / FTM2D)
through the C-ABI on one GPU, synthetic inputs of the SURVEY 8(d) shapes.  Not the bench contract (bench.py is
the Serra09 line); these are the numbers DESIGN.md section 6 quotes for the remaining configs.

    python tools/bench_configs.py [--quick]

Every figure is wall clock around the library call (host buffers in, host results out), after one warm-up call.
Pair counts are bounded samples of the full grids; the per-pair rates extrapolate linearly.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import acoss_b200
from acoss_b200 import synth


def timed(fn, reps=2):
    fn()
    t0 = time.perf_counter()
    for _ in range(reps):
        out = fn()
    return (time.perf_counter() - t0) / reps, out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true")
    args = ap.parse_args()
    q = 0.25 if args.quick else 1.0
    ctx = acoss_b200.Context(0)
    res = {}
    rng = np.random.default_rng(0)

    # ---- C4 LateFusionChen: Qmax + Dmax on the C2-L tracks, SNF of two 1000 x 1000 matrices (K = 20, 20 iterations)
    ds = synth.make_named("datacos1000", "L", scale=0.1)
    tracks = [np.ascontiguousarray(h, np.float32) for h in ds["hpcp"]]
    ctx.upload_tracks(tracks)
    n = len(tracks)
    i, j = np.triu_indices(n, 1)
    pairs = np.stack([i, j], 1).astype(np.int32)
    t, _ = timed(lambda: ctx.serra09_pairs(pairs, want_dmax=True))
    res["C4_chen_qmax_dmax"] = {"pairs": len(pairs), "seconds": t, "pairs_per_s": len(pairs) / t,
                                "note": "%d C2-L tracks (n~U[1600,2400]), CRP built once, Qmax and Dmax from it" % n}
    nt = int(1000 * q)
    A = rng.random((nt, nt)); A = 0.5 * (A + A.T) + 0.1
    B = rng.random((nt, nt)); B = 0.5 * (B + B.T) + 0.1
    if hasattr(ctx, "snf"):
        t, _ = timed(lambda: ctx.snf([A, B], K=20, niters=20, reg_diag=1, want_Ws=False), reps=1)
        res["C4_snf"] = {"n": nt, "n_mats": 2, "K": 20, "niters": 20, "seconds": t}

    # ---- C3 EarlyFusionTralie: block features (300-600 blocks; mfccs 1000-d, ssms 1225-d, chromas 480-d)
    nw = max(2, int(10 * q))
    bf = synth.make_block_features(nw, 2, 4, 1001, 300, 600)
    ctx.earlyfusion_upload(bf["block_feats"])
    n = len(bf["block_feats"])
    i, j = np.triu_indices(n, 1)
    pairs = np.stack([i, j], 1).astype(np.int32)
    t, _ = timed(lambda: ctx.earlyfusion_pairs(pairs), reps=1)
    nb = np.array([f["mfccs"].shape[0] for f in bf["block_feats"]])
    cells = float(np.sum(nb[pairs[:, 0]] * nb[pairs[:, 1]]))
    res["C3_earlyfusion"] = {"pairs": len(pairs), "seconds": t, "pairs_per_s": len(pairs) / t,
                             "gemm_tflops": 2.0 * cells * 2705 / t / 1e12, "sw_gcups": 4.0 * cells / t / 1e9,
                             "note": "%d tracks, 300-600 blocks; 3 CSMs + 4 binarisations + WCSM fusion + 4 constrained SW per pair" % n}

    # ---- C5 SiMPle (ordered pairs) and FTM2D (15 000 x 900 all-pairs)
    ns = int(400 * q)
    feats = []
    for _ in range(ns):
        L = int(rng.integers(150, 251))
        f = rng.random((12, L))
        feats.append(f / np.sqrt(np.sum(f * f, 0, keepdims=True)))
    ctx.simple_upload(feats)
    ii, jj = np.meshgrid(np.arange(ns), np.arange(ns), indexing="ij")
    keep = ii != jj
    pairs = np.stack([ii[keep], jj[keep]], 1).astype(np.int32)
    t, _ = timed(lambda: ctx.simple_pairs(pairs), reps=1)
    res["C5_simple"] = {"ordered_pairs": len(pairs), "seconds": t, "pairs_per_s": len(pairs) / t,
                        "note": "%d tracks, (12, L) float64, L~U[150,250], SSLEN 10" % ns}
    nf = int(15000 * q)
    S = rng.random((nf, 900))
    S /= np.sqrt(np.sum(S * S, 1, keepdims=True))
    t, _ = timed(lambda: ctx.ftm2d_allpairs(S), reps=1)
    res["C5_ftm2d"] = {"tracks": nf, "seconds": t, "pairs_per_s": nf * (nf - 1) / 2 / t,
                       "fp64_tflops": 2.0 * 900 * nf * nf / t / 1e12,
                       "note": "exp(-|s_i - s_j|^2) for all pairs, float64 in and out (N x N matrix to the host inside the timed call)"}
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
