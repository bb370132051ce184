#!/usr/bin/env python
"""Where the host-side time of one end-to-end Serra09 step goes (bench.py's e2e leg under cProfile), C2-L, one GPU."""
import cProfile, pstats, io, os, sys, tempfile, contextlib, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import acoss_b200
from acoss_b200 import synth


def main():
    scale = float(sys.argv[1]) if len(sys.argv) > 1 else 1.0
    ds = synth.make_named("datacos1000", "L", scale=scale, hard=True)
    tracks = ds["hpcp"]
    tmp = tempfile.mkdtemp(prefix="acoss_e2e_")
    os.chdir(tmp)
    csv = os.path.join(tmp, "ds.csv")
    with open(csv, "w") as f:
        f.write("work_id,track_id\n")
        for w, t in zip(ds["work_ids"], ds["track_ids"]):
            f.write("%s,%s\n" % (w, t))
    sink = io.StringIO()
    with contextlib.redirect_stdout(sink):
        s = acoss_b200.Serra09(csv, os.path.join(tmp, "feat") + "/", shortname="bench", downsample_fac=ds["fac"],
                               cachedir=os.path.join(tmp, "cache"), device=0)
    for k, t in enumerate(tracks):
        s.all_feats[k] = t
        s.cliques.setdefault(ds["labels"][k], set([])).add(k)

    def step():
        with contextlib.redirect_stdout(sink):
            s._uploaded = False
            s.Ds["main"][:, :] = 0
            s.all_pairwise(1, symmetric=True)
            s.normalize_by_length()
            return s.getEvalStatistics("main")
    step()
    torch.cuda.synchronize()
    pr = cProfile.Profile()
    t0 = time.perf_counter()
    pr.enable(); step(); torch.cuda.synchronize(); pr.disable()
    print("step %.1f ms; pair kernel time %.1f ms" % ((time.perf_counter() - t0) * 1e3, s.last_stats["ms_total"]))
    out = io.StringIO()
    pstats.Stats(pr, stream=out).sort_stats("cumulative").print_stats(28)
    print(out.getvalue())


main()
