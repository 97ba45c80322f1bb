#!/usr/bin/env python3
"""Summarise `ncu --page source --print-source cuda,sass --csv`: per source line, instructions executed and stall samples."""
import csv, sys
path = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur = None; rows = []
for r in csv.reader(open(path)):
    if len(r) == 2 and r[0] == "File Path": cur = r[1].split("/")[-1]; continue
    if len(r) < 10 or r[0] in ("Line No", ""): continue
    try:
        rows.append((cur, int(r[0]), r[1].strip(), int(r[6]), int(r[7])))
    except ValueError:
        pass
tot_i = sum(x[4] for x in rows); tot_s = sum(x[3] for x in rows)
print(f"total inst {tot_i:.3e} samples {tot_s}")
for f, ln, src, s, i in sorted(rows, key=lambda x: -x[4])[:top]:
    print(f"{100*i/tot_i:5.1f}% inst {100*s/max(tot_s,1):5.1f}% smp  {f}:{ln:<4d} {src[:110]}")
