#!/usr/bin/env python
"""Executed warp-instructions and stall samples of one kernel, bucketed by 20-line source regions.
    python tools/ncu_buckets.py <report.ncu-rep> <library.so> <kernel-name-substring> <units (robots)>"""
import csv
import importlib.util
import os
import subprocess
import sys
from collections import defaultdict

spec = importlib.util.spec_from_file_location("nbl", os.path.join(os.path.dirname(os.path.abspath(__file__)), "ncu_by_line.py"))
nbl = importlib.util.module_from_spec(spec)
spec.loader.exec_module(nbl)


def main():
    rep, so, kernel, units = sys.argv[1], sys.argv[2], sys.argv[3], float(sys.argv[4])
    mangled = sys.argv[5] if len(sys.argv) > 5 else kernel   # substring of the mangled name when `kernel` matches several instantiations
    table = nbl.line_table(so, mangled)
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "-k", f"regex:{kernel}", "-c", "1"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr = rows[1]
    iA, iI, iS = hdr.index("Address"), hdr.index("Instructions Executed"), hdr.index("# Samples")
    b = defaultdict(lambda: [0.0, 0.0])
    base, ti, ts = None, 0.0, 0.0
    for r in rows[2:]:
        if len(r) < len(hdr):
            continue
        a = int(r[iA], 16) if r[iA].startswith("0x") else int(r[iA])
        if base is None:
            base = a
        ie, sm = float(r[iI] or 0), float(r[iS] or 0)
        ti += ie
        ts += sm
        loc = table.get(a - base)
        key = ("unknown", 0) if not loc else (loc[0][0], (loc[0][1] // 20) * 20)
        b[key][0] += ie
        b[key][1] += sm
    print(f"total warp-instructions per unit: {ti / units:.0f}")
    for (f, l), (ie, sm) in sorted(b.items()):
        if ie / ti > 0.008 or sm / ts > 0.008:
            print(f"{f}:{l:<4d}-{l + 19:<4d} inst/unit {ie / units:7.0f} ({100 * ie / ti:4.1f}%)  stall samples {100 * sm / ts:4.1f}%")


if __name__ == "__main__":
    main()
