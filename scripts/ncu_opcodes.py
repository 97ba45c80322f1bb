#!/usr/bin/env python3
"""Executed-instruction histogram by SASS opcode from `ncu --page source --print-source cuda,sass --csv` (each SASS row
carries its executed warp-instruction count).  Usage: ncu_opcodes.py src.csv [top]"""
import csv, re, sys, collections
path = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
ops = collections.Counter(); seen = set(); total = 0
PIPE = {"FFMA": "fma", "FMUL": "fma", "FADD": "fma", "FSETP": "alu/fp", "FSEL": "alu", "FMNMX": "alu", "FMNMX3": "alu"}
for r in csv.reader(open(path)):
    if len(r) < 10 or r[0] != "" or not r[2].startswith("0x"): continue
    addr = r[2]
    if addr in seen: continue            # a SASS instruction is listed once per source line it is attributed to
    seen.add(addr)
    try: n = int(r[7])
    except ValueError: continue
    m = re.match(r"\s*(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", r[3])
    if not m: continue
    ops[m.group(1)] += n; total += n
print(f"executed warp instructions {total:.4e} over {len(seen)} SASS instructions")
acc = 0
for k, v in ops.most_common(top):
    acc += v
    print(f"{100*v/total:6.2f}%  {100*acc/total:6.2f}%  {k}")
