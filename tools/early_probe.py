#!/usr/bin/env python
"""Ad-hoc: EarlyFusionTralie pair pipeline on 24 synthetic tracks (276 pairs, 300-600 blocks), wall clock."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import acoss_b200
from acoss_b200 import synth
ctx = acoss_b200.Context(0)
bf = synth.make_block_features(10, 2, 4, 1001, 300, 600)
ctx.earlyfusion_upload(bf["block_feats"])
n = len(bf["block_feats"])
i, j = np.triu_indices(n, 1)
pairs = np.stack([i, j], 1).astype(np.int32)
ctx.earlyfusion_pairs(pairs)
t0 = time.perf_counter()
out = ctx.earlyfusion_pairs(pairs)
dt = time.perf_counter() - t0
print("pairs %d  %.4f s  %.0f pairs/s  checksum %.9f" % (len(pairs), dt, len(pairs) / dt, float(np.sum(out))))
