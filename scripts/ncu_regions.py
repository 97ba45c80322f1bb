#!/usr/bin/env python3
"""Executed warp instructions of one kernel by REGION of the source (file + line ranges), with the share of FP32 math,
moves, local-memory traffic and control flow inside each region.  Input: ncu --page source --print-source cuda,sass --csv.
Usage: ncu_regions.py src.csv"""
import csv, re, sys, collections
REGIONS = [   # line ranges of the round-2 final sources (rt_kernels.cuh 729, rt_trace.cuh 597, rt_device.cuh 632 lines)
    ("tile prologue: MT19937 twists, checkpoint load, tile mask box", [("rt_kernels.cuh", 229, 252), ("mt19937.cuh", 1, 999), ("rt_kernels.cuh", 184, 211)]),
    ("round: sample set-up, jitter, camera ray (FP64)", [("rt_kernels.cuh", 253, 297), ("rt_kernels.cuh", 114, 128)]),
    ("round: resolve + store", [("rt_kernels.cuh", 298, 320), ("rt_kernels.cuh", 130, 160)]),
    ("round masks: box reduce + cand_build + bundle tests", [("rt_device.cuh", 185, 323), ("rt_cull.cuh", 1, 999), ("rt_kernels.cuh", 162, 172), ("device_atomic_functions.hpp", 1, 9999)]),
    ("scan loops (masked / run-based) + scan glue", [("rt_trace.cuh", 68, 280)]),
    ("grid walk (rt_grid.cuh)", [("rt_grid.cuh", 1, 999)]),
    ("sphere_t / plane_t / sphere_disc / sphere_mask", [("rt_device.cuh", 324, 427)]),
    ("generic_hit + run_program + leaf_interval + csg", [("rt_trace.cuh", 15, 67), ("rt_device.cuh", 428, 632), ("rt_csg.cuh", 1, 999)]),
    ("closest_hit finalise (normal, material, code)", [("rt_trace.cuh", 281, 330)]),
    ("shade loop + light_vector + light_shade", [("rt_trace.cuh", 331, 478)]),
    ("ray_step / trace (recursion glue)", [("rt_trace.cuh", 479, 597)]),
    ("wavefront kernel body (queue, state)", [("rt_kernels.cuh", 321, 528)]),
    ("vector helpers (dot, madd, normalized, SV accessors)", [("rt_device.cuh", 60, 184)]),
]
CLASS = {"fp32": {"FFMA", "FMUL", "FADD", "FSETP", "FSEL", "FMNMX", "FMNMX3", "MUFU", "FCHK", "FFMA2", "FMUL2", "FADD2"},
         "fp64": {"DFMA", "DMUL", "DADD", "DSETP", "F2F", "I2F", "F2I", "DMNMX"},
         "move": {"MOV", "IMAD.MOV", "PRMT", "SEL", "CS2R", "UMOV", "R2UR", "S2R", "S2UR", "MOVM", "SHFL", "R2P", "P2R", "BMOV"},
         "local": {"LDL", "STL"},
         "mem": {"LDS", "STS", "LDG", "STG", "LDC", "LDCU", "ATOMS", "ATOMG", "RED", "LD", "ST"},
         "ctrl": {"BRA", "BSSY", "BSYNC", "BREAK", "CALL", "RET", "EXIT", "BAR", "WARPSYNC", "NANOSLEEP", "YIELD", "JMP", "BRX"},
         }
def klass(op, full):
    if op == "IMAD" and (".MOV" in full or re.search(r"IMAD(\.\w+)* R\d+, RZ, RZ,", full)): return "move"
    for k, s in CLASS.items():
        if op in s: return k
    return "int"
path = sys.argv[1]
cur = None; seen = set()
reg_tot = collections.Counter(); reg_cls = collections.defaultdict(collections.Counter); total = 0
def region_of(f, ln):
    for name, spans in REGIONS:
        for ff, a, b in spans:
            if f == ff and a <= ln <= b: return name
    return "other: " + str(f)
line_of_sass = {}
rows = list(csv.reader(open(path)))
# pass 1: source lines carry their SASS underneath?  The csv lists per source line the aggregated counts, then (with sass) per
# SASS row: we need file+line per SASS instruction: rows with r[0]=='' are SASS rows following their source line row.
f = None; ln = None
for r in rows:
    if len(r) == 2 and r[0] == "File Path": f = r[1].split("/")[-1]; continue
    if len(r) < 10: continue
    if r[0] not in ("", "Line No"):
        try: ln = int(r[0])
        except ValueError: pass
        continue
    if r[0] == "" and r[2].startswith("0x"):
        addr = r[2]
        if addr in seen: continue
        seen.add(addr)
        try: n = int(r[7])
        except ValueError: continue
        m = re.match(r"\s*(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", r[3])
        if not m: continue
        op = m.group(1); reg = region_of(f, ln)
        reg_tot[reg] += n; reg_cls[reg][klass(op, r[3])] += n; total += n
print(f"executed warp instructions {total:.4e}")
print(f"{'region':62s} {'share':>6s} | " + " ".join(f"{k:>6s}" for k in ("fp32", "fp64", "int", "move", "local", "mem", "ctrl")))
for reg, n in reg_tot.most_common():
    c = reg_cls[reg]
    print(f"{reg:62s} {100*n/total:5.1f}% | " + " ".join(f"{100*c[k]/total:5.1f}%" for k in ("fp32", "fp64", "int", "move", "local", "mem", "ctrl")))
allc = collections.Counter()
for c in reg_cls.values(): allc.update(c)
print(f"{'all':62s} 100.0% | " + " ".join(f"{100*allc[k]/total:5.1f}%" for k in ("fp32", "fp64", "int", "move", "local", "mem", "ctrl")))
