#!/usr/bin/env python
"""profiles/traffic.json from an ncu launch list of tools/crp_probe.py:

    ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv \\
        --log-file launches.csv python tools/crp_probe.py <L> <mode>
    python tools/ncu_traffic_launches.py launches.csv <L> <n_tracks> <calls> "<note>" > profiles/traffic.json

DRAM bytes (read + write) of all CRP-stage launches per CRP cell, of the DP launches per CRP byte (M x pitch, pitch = 128-byte
multiple), and the per-kernel table the README quotes."""
import collections
import csv
import json
import re
import sys


def main(path, L, n_tracks, calls, note):
    lines = [l for l in open(path, errors="replace") if not l.startswith("==")]
    acc = collections.OrderedDict()
    for x in csv.DictReader(lines):
        name = re.sub(r"\(.*", "", x["Kernel Name"]).replace("void ", "")
        v = float(x["Metric Value"].replace(",", ""))
        u = x["Metric Unit"]
        a = acc.setdefault(name, {"launches": 0, "ms": 0.0, "dram_bytes": 0.0})
        if x["Metric Name"] == "gpu__time_duration.sum":
            a["ms"] += v * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}[u]
            a["launches"] += 1
        elif x["Metric Name"].startswith("dram__bytes"):
            a["dram_bytes"] += v * {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]
    M = L - 8
    pairs = n_tracks * (n_tracks - 1) // 2
    cells = float(calls) * pairs * M * M
    pitch = (M + 127) // 128 * 128
    crp_bytes = float(calls) * pairs * M * pitch
    crp = sum(a["dram_bytes"] for k, a in acc.items() if k.startswith("crp_"))
    dp = sum(a["dram_bytes"] for k, a in acc.items() if k.startswith("dp_"))
    out = {"source": note, "captured_launches": {"crp": sum(a["launches"] for k, a in acc.items() if k.startswith("crp_")),
                                                  "dp": sum(a["launches"] for k, a in acc.items() if k.startswith("dp_"))},
           "crp_dram_bytes_per_cell": crp / cells, "crp_dram_bytes_captured": crp,
           "dp_dram_bytes_per_crp_byte": dp / crp_bytes, "dp_dram_bytes_captured": dp,
           "per_kernel": {k: a for k, a in acc.items() if a["ms"] > 0.01}}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), sys.argv[5])
