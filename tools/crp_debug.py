#!/usr/bin/env python
"""Ad-hoc: compare one CRP mode against mode 1 on a few pairs and print where they differ."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import acoss_b200
from acoss_b200 import synth

def main():
    mode = int(sys.argv[1])
    lens = [int(x) for x in sys.argv[2].split(",")]
    ctx = acoss_b200.Context(0)
    labels, works, tids, hp = synth.make_tracks(len(lens), 1, 0, 5, max(lens), max(lens))
    hp = [h[:L] for h, L in zip(hp, lens)]
    ctx.upload_tracks(hp)
    for a in range(len(lens)):
        b = (a + 1) % len(lens)
        R1, q1, r1, o1 = ctx.crp_pair(a, b, ctx.params(crp_mode=1))
        R2, q2, r2, o2 = ctx.crp_pair(a, b, ctx.params(crp_mode=mode))
        c = ctx.debug_counters()
        st = ctx.pair_stats()
        print("pair", a, b, "shape", R1.shape, "oti", o1, o2, "thr_q diff", int((q1 != q2).sum()), "thr_r diff", int((r1 != r2).sum()),
              "R diff", int((R1 != R2).sum()), "ones", int(R1.sum()), int(R2.sum()),
              "exact", st["n_exact_cells"], "fallback", st["n_fallback_rows"],
              "maxerr/2^-24", np.array([c[14]], np.uint32).view(np.float32)[0] * 2 ** 24 if len(c) > 14 else None)
        if (r1 != r2).any():
            w = np.nonzero(r1 != r2)[0]
            print("  thr_r first diffs", w[:8], r1[w[:4]], r2[w[:4]])
        if (q1 != q2).any():
            w = np.nonzero(q1 != q2)[0]
            print("  thr_q first diffs", w[:8], q1[w[:4]], q2[w[:4]])
        if (R1 != R2).any():
            w = np.argwhere(R1 != R2)
            print("  R first diffs", w[:8].tolist())
    print("counters", list(c))

main()
