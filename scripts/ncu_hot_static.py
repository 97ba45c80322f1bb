#!/usr/bin/env python3
"""Hot static footprint by source line from `ncu --page source --print-source cuda,sass --csv`."""
import csv, sys
path = sys.argv[1]; thr = float(sys.argv[2]) if len(sys.argv) > 2 else 2e6; top = int(sys.argv[3]) if len(sys.argv) > 3 else 60
cur = None; line = None; hot = {}; tot = 0
for r in csv.reader(open(path)):
    if len(r) == 2 and r[0] == "File Path": cur = r[1].split("/")[-1]; continue
    if len(r) < 10 or r[0] == "Line No": continue
    if r[0] != "":
        try: line = (cur, int(r[0]), r[1].strip()[:90])
        except ValueError: line = None
        continue
    try: e = int(r[7])
    except ValueError: continue
    if e >= thr:
        hot.setdefault(line, [0, 0]); hot[line][0] += 1; hot[line][1] += e; tot += 1
print("hot static instrs:", tot)
for k, v in sorted(hot.items(), key=lambda x: -x[1][0])[:top]:
    print(f"{v[0]:4d} static  {v[1]/1e9:7.3f}G exec  {k[0]}:{k[1]}  {k[2]}")
