#!/usr/bin/env python
"""profiles/traffic.json from `ncu -i X.ncu-rep --page raw --csv`: DRAM bytes (read + write) of the captured
crp_strip_gemm_kernel launches per CRP cell, and of the dp_rowpar_kernel launches per CRP byte.

    python tools/ncu_traffic.py raw.csv <crp_cells_in_the_captured_launches> <crp_bytes_read_by_the_captured_dp_launches> "<source note>"
"""
import csv
import json
import sys


def main(path, crp_cells, dp_bytes, note):
    rows = list(csv.reader(open(path, errors="replace")))
    hdr = rows[0]
    i_name = hdr.index("Kernel Name")
    i_rd = hdr.index("dram__bytes_read.sum")
    i_wr = hdr.index("dram__bytes_write.sum")
    units = rows[1]
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    acc = {"crp": 0.0, "dp": 0.0}
    n = {"crp": 0, "dp": 0}
    for r in rows[2:]:
        if len(r) <= max(i_rd, i_wr):
            continue
        b = float(r[i_rd]) * scale[units[i_rd]] + float(r[i_wr]) * scale[units[i_wr]]
        if "crp_strip_gemm" in r[i_name]:
            acc["crp"] += b; n["crp"] += 1
        elif "dp_rowpar" in r[i_name]:
            acc["dp"] += b; n["dp"] += 1
    out = {"source": note, "captured_launches": n}
    if n["crp"]:
        out["crp_dram_bytes_per_cell"] = acc["crp"] / float(crp_cells)
        out["crp_dram_bytes_captured"] = acc["crp"]
    if n["dp"]:
        out["dp_dram_bytes_per_crp_byte"] = acc["dp"] / float(dp_bytes)
        out["dp_dram_bytes_captured"] = acc["dp"]
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main(sys.argv[1], float(sys.argv[2]), float(sys.argv[3]), sys.argv[4])
