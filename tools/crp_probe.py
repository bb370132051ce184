#!/usr/bin/env python
"""Ad-hoc probe: CRP-stage throughput (cells/s) of the pair kernel at fixed track lengths."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import acoss_b200
from acoss_b200 import synth

def main():
    lens = [int(x) for x in sys.argv[1].split(",")]
    mode = int(sys.argv[2]) if len(sys.argv) > 2 else 2
    n = int(os.environ.get("ACOSS_PROBE_N", "64"))
    ctx = acoss_b200.Context(0)
    for kv in os.environ.get("ACOSS_PROBE_OPTS", "").split(","):
        if kv:
            k, v = kv.split("=")
            ctx.set_option(k, int(v))
    for L in lens:
        labels, works, tids, hp = synth.make_tracks(n // 2, 2, 0, 5, L, L)
        hp = [h[:L] for h in hp]
        if os.environ.get("ACOSS_PROBE_IID"):        # frames without any temporal structure (the worst case for the spread sample)
            rng = np.random.default_rng(7)
            base = rng.random((L + 50, 12)).astype(np.float32)
            hp = []
            for k in range(n):
                x = np.roll(base[k % 7: k % 7 + L], k % 12, axis=1) + 0.08 * rng.random((L, 12)).astype(np.float32)
                hp.append(np.ascontiguousarray(x / x.max(axis=1, keepdims=True), np.float32))
        ctx.upload_tracks(hp)
        i, j = np.triu_indices(n, 1)
        pairs = np.stack([i, j], 1).astype(np.int32)
        prm = ctx.params(crp_mode=mode)
        ctx.serra09_pairs(pairs, prm)
        c0 = ctx.debug_counters_ext(16) if hasattr(ctx, "debug_counters_ext") else None
        ctx.serra09_pairs(pairs, prm)
        st = ctx.pair_stats()
        if c0 is not None:
            c1 = ctx.debug_counters_ext(16)
            import struct
            err = struct.unpack("f", struct.pack("I", int(c1[14]) & 0xFFFFFFFF))[0]
            print("  counters", [int(b) for b in c1], " max |rho~ - rho| / (xx + yy) on the exact cells = %.1f u (bound 2048 u)" % (err * 2.0 ** 24))
        print("L=%d mode=%d pairs=%d crp %.1f ms  %.3e cells/s   dp %.2f ms %.0f GCUPS  exact %.4f fallback %d" % (
            L, mode, len(pairs), st["ms_crp"], st["crp_cells"] / st["ms_crp"] * 1e3, st["ms_dp"],
            st["dp_cells"] / st["ms_dp"] / 1e6, st["n_exact_cells"] / (2.0 * st["crp_cells"]), st["n_fallback_rows"]))

main()
