#!/usr/bin/env python
"""Per-source-line summary of an `ncu --page source --csv --print-source cuda,sass` export:
instructions executed and stall samples per CUDA source line (top N)."""
import csv
import sys


def main(path, top=40):
    rows = list(csv.reader(open(path, errors="replace")))
    cur_file = None
    hdr = None
    agg = {}
    for r in rows:
        if not r:
            continue
        if r[0] == "File Name":
            cur_file = r[1].split("/")[-1]
            continue
        if r[0] == "Line No":
            hdr = r
            i_inst = hdr.index("Instructions Executed")
            i_samp = hdr.index("# Samples")
            continue
        if hdr is None or len(r) <= i_inst:
            continue
        if r[0] != "":            # a CUDA source line row (aggregated over its SASS)
            try:
                inst = int(r[i_inst]); samp = int(r[i_samp])
            except ValueError:
                continue
            agg[(cur_file, int(r[0]))] = (inst, samp, r[1].strip()[:100])
    tot_i = sum(v[0] for v in agg.values()) or 1
    tot_s = sum(v[1] for v in agg.values()) or 1
    print("total warp-instructions %d, samples %d" % (tot_i, tot_s))
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
        print("%-22s %4d  inst %5.1f%%  samples %5.1f%%  %s" % (k[0], k[1], 100.0 * v[0] / tot_i, 100.0 * v[1] / tot_s, v[2]))


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40)
