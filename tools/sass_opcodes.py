#!/usr/bin/env python
"""Per-kernel counts of the SASS opcodes that prove (or disprove) what a kernel claims to use: tcgen05 MMA (UTC*MMA), TMEM
loads / stores (LDTM / STTM), bulk and tensor TMA copies (UBLKCP / UTMALDG), DPX (VIMNMX3 / VIADDMNMX), FP64 FMA (DFMA),
plus FFMA / HFMA2 for scale.  Reads `cuobjdump -sass` of the built library and writes profiles/sass_opcodes.txt.

    python tools/sass_opcodes.py [path/to/lib.so] [out.txt]
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OPS = ["UTCHMMA", "UTCQMMA", "UTCIMMA", "UTCOMMA", "LDTM", "STTM", "UTCBAR", "UBLKCP", "UTMALDG", "UTMASTG", "SYNCS",
       "VIMNMX3", "VIADDMNMX", "VIMNMX", "DFMA", "DMMA", "HMMA", "FFMA", "HFMA2", "LDG", "STG", "LDS", "STS", "ATOMS", "RED", "ATOMG", "REDUX"]


def main():
    lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "acoss-1_b200", "csrc", "libacoss_b200.so")
    out = sys.argv[2] if len(sys.argv) > 2 else os.path.join(ROOT, "profiles", "sass_opcodes.txt")
    txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
    counts = collections.OrderedDict()
    cur = None
    for line in txt.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip() or m.group(1)
            cur = re.sub(r"\(.*", "", cur)
            counts[cur] = collections.Counter()
            continue
        if cur is None:
            continue
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
        if m:
            op = m.group(1)
            counts[cur]["_total"] += 1
            for o in OPS:
                if op == o or op.startswith(o + "."):
                    counts[cur][o] += 1
    with open(out, "w") as f:
        f.write("# %s\n# per-kernel SASS opcode counts (cuobjdump -sass, sm_100a); 0 entries omitted\n" % os.path.relpath(lib, ROOT))
        for k, c in counts.items():
            items = " ".join("%s=%d" % (o, c[o]) for o in OPS if c[o])
            f.write("%-60s total=%-6d %s\n" % (k[:60], c["_total"], items))
    print("wrote", out)


if __name__ == "__main__":
    main()
