#!/usr/bin/env python
"""Ad-hoc: EarlyFusionTralie pair pipeline on 24 synthetic tracks (276 pairs, 300-600 blocks), wall clock; the cluster
(TMA-multicast) variant of the CSM GEMM against the one-CTA-per-tile variant: scores must be bit-identical."""
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
ref = None
for cl in (0, 1):
    ctx.set_option("ef_cluster", cl)
    ctx.earlyfusion_pairs(pairs)
    t0 = time.perf_counter()
    out = ctx.earlyfusion_pairs(pairs)
    dt = time.perf_counter() - t0
    ms3 = {k: round(v, 2) for k, v in ctx.earlyfusion_stats().items() if k.startswith("ms_")}
    print("cluster %d: pairs %d  %.4f s  %.0f pairs/s  checksum %.9f  stage ms %s" % (cl, len(pairs), dt, len(pairs) / dt, float(np.sum(out)), ms3))
    if ref is None:
        ref = out
    else:
        print("identical to the one-CTA variant:", bool(np.array_equal(ref, out)))
