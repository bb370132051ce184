#!/usr/bin/env python
"""Ad-hoc: failure-reason counters of a CRP mode on the large parity-test tracks."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import acoss_b200
from test_gpu_parity import _tracks

mode = int(sys.argv[1])
noise = float(sys.argv[2]) if len(sys.argv) > 2 else 0.08
ctx = acoss_b200.Context(0)
rng = np.random.default_rng(33)
tracks = _tracks(rng, [int(x) for x in rng.integers(1600, 2400, size=10)], noise=noise)
ctx.upload_tracks(tracks)
pairs = np.array([(i, j) for i in range(10) for j in range(10) if i < j], np.int32)
q2 = ctx.serra09_pairs(pairs, ctx.params(crp_mode=mode))
st = ctx.pair_stats()
c = ctx.debug_counters()
print("mode", mode, "exact frac", st["n_exact_cells"] / (2.0 * st["crp_cells"]), "fallback rows", st["n_fallback_rows"],
      "of", int(sum(t.shape[0] - 8 for t in tracks) * 9), "ms_crp", st["ms_crp"])
print("counters", list(c), "maxerr u", np.array([c[14]], np.uint32).view(np.float32)[0] * 2 ** 24)
