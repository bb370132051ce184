#!/usr/bin/env python
"""Ad-hoc: per-phase cycle counters of the mode-4 CRP kernel at one track length."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import acoss_b200
from acoss_b200 import synth

L = int(sys.argv[1]); mode = int(sys.argv[2])
n = 64
ctx = acoss_b200.Context(0)
labels, works, tids, hp = synth.make_tracks(n // 2, 2, 0, 5, L, L)
hp = [h[:L] for h in hp]
ctx.upload_tracks(hp)
i, j = np.triu_indices(n, 1)
pairs = np.stack([i, j], 1).astype(np.int32)
prm = ctx.params(crp_mode=mode)
ctx.serra09_pairs(pairs, prm)
ctx.serra09_pairs(pairs, prm)
st = ctx.pair_stats(); c = ctx.debug_counters()
tasks = max(1, c[7]); tiles = (L - 8 + 255) // 256
print("dbg=%s L=%d crp %.1f ms; per strip kclk: sample %.1f pivots %.1f count %.1f (%.2f/tile) select %.1f binarise(pass2 strips) %.1f (%.2f/tile); fallback %d ovf %d miss %d" % (
    os.environ.get("ACOSS_GEMM_DBG"), L, st["ms_crp"], c[2] / tasks / 1e3, c[3] / tasks / 1e3, c[4] / tasks / 1e3, c[4] / tasks / 1e3 / tiles,
    c[5] / tasks / 1e3, c[6] / (tasks / 2) / 1e3, c[6] / (tasks / 2) / 1e3 / tiles, c[1], c[15], c[13]))
x = ctx.debug_counters_ext(48)[16:]
names = ["stageA+ymax", "smp gather", "smp mma wait", "piv dump", "piv hist", "count pass", "sel list load", "sel binning", "sel scan",
         "sel collect", "sel bisect rows", "sel offsets", "sel exact chains", "sel rank", "fallback rows", "bin stream", "bin exact"]
print("sub-phases (kclk per strip): " + "; ".join("%s %.1f" % (nm, v / tasks / 1e3) for nm, v in zip(names, x)))
print("stats: crp_cells %d crp_bytes %d dp_cells %d ms_dp %.3f n_exact %d" % (st["crp_cells"], st["crp_bytes"], st["dp_cells"], st["ms_dp"], st["n_exact_cells"]))
