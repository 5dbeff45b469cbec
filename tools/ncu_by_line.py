#!/usr/bin/env python
"""Attribute the per-SASS-instruction counters of an .ncu-rep to CUDA source lines.

    python tools/ncu_by_line.py <report.ncu-rep> <library.so> <kernel-name-substring> [top]

ncu's CSV source page only carries SASS rows; the line table comes from `nvdisasm -g` on the cubin."""
import csv
import re
import subprocess
import sys
import tempfile
import os
from collections import defaultdict


def line_table(so, kernel):
    d = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=d, capture_output=True)
    txt = []   # the library is built from several translation units: one cubin each
    for cubin in sorted(f for f in os.listdir(d) if f.endswith(".cubin")):
        txt += subprocess.run(["nvdisasm", "-g", "-c", os.path.join(d, cubin)], capture_output=True, text=True).stdout.splitlines()
    table, cur, inside = {}, None, False
    for ln in txt:
        if ln.startswith(".text."):
            inside = kernel in ln
            continue
        if not inside:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)(?: inlined at "([^"]+)", line (\d+))?', ln)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)), os.path.basename(m.group(3)) if m.group(3) else None, int(m.group(4)) if m.group(4) else None)
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
        if m and cur:
            table[int(m.group(1), 16)] = (cur, m.group(2))
    return table


def main():
    rep, so, kernel = sys.argv[1:4]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr = rows[1]
    iA, iI, iS = hdr.index("Address"), hdr.index("Instructions Executed"), hdr.index("# Samples")
    table = line_table(so, kernel)
    base = None
    per_line, per_outer = defaultdict(lambda: [0.0, 0.0]), defaultdict(lambda: [0.0, 0.0])
    tot_i = tot_s = 0.0
    for r in rows[2:]:
        if len(r) < len(hdr):
            continue
        a = int(r[iA], 16) if r[iA].startswith("0x") else int(r[iA])
        if base is None:
            base = a
        ie, sm = float(r[iI] or 0), float(r[iS] or 0)
        tot_i += ie
        tot_s += sm
        loc = table.get(a - base)
        if not loc:
            continue
        (f, l, of, ol), _ = loc
        per_line[(f, l)][0] += ie
        per_line[(f, l)][1] += sm
        key = (of, ol) if of else (f, l)
        per_outer[key][0] += ie
        per_outer[key][1] += sm
    print(f"total warp-instructions {tot_i:.3e}, samples {tot_s:.0f}")
    for title, d in (("by source line", per_line), ("by outermost (inlined-at) line", per_outer)):
        print(f"--- {title}: top {top} by stall samples")
        for (f, l), (ie, sm) in sorted(d.items(), key=lambda kv: -kv[1][1])[:top]:
            print(f"{f}:{l:<5d} inst {100 * ie / tot_i:5.1f}%  samples {100 * sm / tot_s:5.1f}%")


if __name__ == "__main__":
    main()
