#!/usr/bin/env python
"""Static code size of one kernel by source-line bucket (nvdisasm -g listing on stdin or as argv[1]; argv[2] = kernel
substring).  Complements ncu_buckets.py (dynamic counts): a bucket that is large here but rarely executed is
instruction-cache ballast."""
import collections
import re
import sys

txt = open(sys.argv[1]).read().splitlines()
key = sys.argv[2]
width = int(sys.argv[3]) if len(sys.argv) > 3 else 20
inside, cur = False, ("?", 0)
hist = collections.Counter()
total = 0
for ln in txt:
    if ln.startswith("//---") and ".text." in ln:
        inside = key in ln
        continue
    if not inside:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", ln):
        hist[(cur[0], cur[1] // width * width)] += 1
        total += 1
print(f"{key}: {total} instructions = {total * 16 / 1024:.1f} KiB")
for (f, b), c in sorted(hist.items()):
    if c >= total * 0.004:
        print(f"{f}:{b:<4d}-{b + width - 1:<4d} {c:6d} ({100.0 * c / total:4.1f}%)")
