#!/usr/bin/env python3
"""Static SASS size per source line of one kernel: nvdisasm -g <cubin> > all.dis; sass_static.py all.dis <kernel-substr>"""
import re, sys
from collections import Counter
path, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
cnt = Counter(); fcnt = Counter(); cur = None; inside = False; total = 0
for line in open(path, errors="replace"):
    if line.startswith("//---") and ".text." in line:
        inside = kern in line; continue
    if not inside: continue
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)(?: inlined at "([^"]+)", line (\d+))?', line)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2))); continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s", line):
        cnt[cur] += 1; total += 1
        if cur: fcnt[cur[0]] += 1
print("total", total)
for f, c in fcnt.most_common(): print(f"  {c:6d} {f}")
for (k, c) in cnt.most_common(top): print(f"{c:6d} {k}")
