#!/usr/bin/env python3
"""Summarise `ncu --page source --print-source cuda,sass --csv` per SASS function and per enclosing source function
(the source function is found by scanning the .cuh files for definitions; inlined helpers count where they are written)."""
import csv, os, re, sys, collections
path = sys.argv[1]
src_dir = sys.argv[2] if len(sys.argv) > 2 else os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "raytracing-project_b200", "csrc")
defs = {}
for fn in os.listdir(src_dir):
    if not fn.endswith((".cuh", ".cu")): continue
    marks = []
    for i, line in enumerate(open(os.path.join(src_dir, fn)), 1):
        m = re.match(r"^\s*(?:template\s*<[^>]*>\s*)?(?:static\s+)?(?:__device__|__global__|__host__)[^;(]*?\b([A-Za-z_]\w*)\s*\(", line)
        if m and not line.strip().startswith("//"): marks.append((i, m.group(1)))
    defs[fn] = marks
def enclosing(fn, ln):
    name = "?"
    for i, n in defs.get(fn, []):
        if i <= ln: name = n
        else: break
    return name
cur_file = cur_fn = None
per_sass = collections.Counter(); per_src = collections.Counter(); smp_src = collections.Counter(); smp_sass = collections.Counter()
for r in csv.reader(open(path)):
    if len(r) == 2 and r[0] == "File Path": cur_file = r[1].split("/")[-1]; continue
    if len(r) == 2 and r[0] == "Function Name": cur_fn = re.sub(r"\(.*", "", r[1]).replace("void ", "").replace("rtd::", ""); continue
    if len(r) < 10 or r[0] in ("Line No", ""): continue
    try: ln, s, i = int(r[0]), int(r[6]), int(r[7])
    except ValueError: continue
    per_sass[cur_fn] += i; smp_sass[cur_fn] += s
    k = (cur_fn.split("<")[0], enclosing(cur_file, ln)); per_src[k] += i; smp_src[k] += s
ti = sum(per_sass.values()); ts = sum(smp_sass.values())
print(f"total warp-instructions {ti:.3e}, stall samples {ts}")
print("-- per SASS function")
for k, v in per_sass.most_common(): print(f"{100*v/ti:5.1f}% inst {100*smp_sass[k]/ts:5.1f}% smp  {k}")
print("-- per (SASS function, source function the line is written in)")
for k, v in per_src.most_common(40): print(f"{100*v/ti:5.1f}% inst {100*smp_src[k]/ts:5.1f}% smp  {k[0]:28s} {k[1]}")
