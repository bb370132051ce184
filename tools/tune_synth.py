#!/usr/bin/env python
"""Ad-hoc: retrieval quality (MR/MRR/MDR/MAP of Serra09 Qmax) of the synthetic cover model at a given shape / regime for a list
of (noise, p_sub) settings -- used to place synth.HARD so that the C2 benchmark set is not saturated (MAP in 0.4-0.9)."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import acoss_b200
from acoss_b200 import synth
import torch


def main():
    shape, regime, scale = sys.argv[1], sys.argv[2], float(sys.argv[3])
    settings = [tuple(float(x) for x in a.split(",")) for a in sys.argv[4:]]
    ctx = acoss_b200.Context(0)
    dev = torch.device("cuda", 0)
    for noise, p_sub in settings:
        ds = synth.make_named(shape, regime, scale=scale, noise=noise, p_sub=p_sub)
        tracks = ctx.median_downsample(ds["hpcp"], ds["fac"]) if ds["fac"] > 1 else ds["hpcp"]
        n = len(tracks)
        ctx.upload_tracks(tracks)
        i, j = np.triu_indices(n, 1)
        pairs = np.stack([i, j], 1).astype(np.int32)
        t0 = time.time()
        q, _ = ctx.serra09_pairs(pairs)
        dt = time.time() - t0
        D = np.zeros((n, n), np.float32)
        D[i, j] = q
        D += D.T
        D = (D / np.sqrt(np.array([t.shape[0] for t in tracks], np.float64))[None, :]).astype(np.float32)
        cl = {}
        for k, l in enumerate(ds["labels"]):
            cl.setdefault(l, []).append(k)
        cliques = sorted(cl.values(), key=lambda c: -len(c))
        Ks = np.array([len(c) for c in cliques])
        order = np.concatenate(cliques).astype(np.int32)
        csize = np.repeat(Ks, Ks).astype(np.int32)
        cstart = np.repeat(np.concatenate([[0], np.cumsum(Ks)[:-1]]), Ks).astype(np.int32)
        Dd = torch.from_numpy(D).to(dev)
        out, _ = ctx.eval_statistics(Dd.data_ptr(), n, order, csize, cstart, np.array([1, 10, 100, 1000], np.int32))
        print("noise %.2f p_sub %.2f: %d tracks %.1f s  MR %.3f MRR %.3f MDR %.1f MAP %.4f top1 %d" % (
            noise, p_sub, n, dt, out[0], out[1], out[2], out[3], out[4]), flush=True)


main()
