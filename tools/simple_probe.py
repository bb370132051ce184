#!/usr/bin/env python
"""Ad-hoc: SiMPle pair kernel on 120 synthetic tracks (14 280 ordered pairs), wall clock."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import acoss_b200
ctx = acoss_b200.Context(0)
rng = np.random.default_rng(0)
ns = int(sys.argv[1]) if len(sys.argv) > 1 else 120
feats = []
for _ in range(ns):
    L = int(rng.integers(150, 251))
    f = rng.random((12, L))
    feats.append(f / np.sqrt(np.sum(f * f, 0, keepdims=True)))
ctx.simple_upload(feats)
ii, jj = np.meshgrid(np.arange(ns), np.arange(ns), indexing="ij")
keep = ii != jj
pairs = np.stack([ii[keep], jj[keep]], 1).astype(np.int32)
ctx.simple_pairs(pairs)
t0 = time.perf_counter()
out = ctx.simple_pairs(pairs)
dt = time.perf_counter() - t0
print("pairs %d  %.4f s  %.0f pairs/s  checksum %.12f" % (len(pairs), dt, len(pairs) / dt, float(np.sum(out))))
