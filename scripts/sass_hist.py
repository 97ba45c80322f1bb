#!/usr/bin/env python3
"""Static SASS opcode histogram of one kernel, split by the out-of-line device functions inside it (nvdisasm -g output).
Usage: sass_hist.py all.dis <kernel-mangled-substring>   (cuobjdump -xelf all rt_api.o; nvdisasm -g *.cubin > all.dis)"""
import collections, re, sys
path, kern = sys.argv[1], sys.argv[2]
GROUPS = [("FP32 math", {"FFMA", "FMUL", "FADD", "FSETP", "FSEL", "FMNMX", "MUFU", "FCHK"}),
          ("FP64", {"DFMA", "DMUL", "DADD", "DSETP", "F2F", "I2F", "F2I", "DMNMX"}),
          ("integer / logic / predicate", {"IMAD", "IADD3", "VIADD", "LOP3", "SHF", "LEA", "ISETP", "PLOP3", "SEL", "FLO", "BREV", "POPC", "PRMT", "VIMNMX", "IABS", "VOTE", "R2P", "P2R", "ULOP3", "UIADD3", "ULEA", "UISETP", "USEL", "UIMAD", "USHF", "UFLO"}),
          ("moves", {"MOV", "UMOV", "R2UR", "S2R", "S2UR", "CS2R", "SHFL", "HFMA2", "BMOV"}),
          ("local memory", {"LDL", "STL"}),
          ("shared / global / constant", {"LDS", "STS", "LDG", "STG", "LDC", "LDCU", "ATOMS", "ATOMG", "RED", "LD", "ST", "CREDUX", "REDUX"}),
          ("control flow", {"BRA", "BSSY", "BSYNC", "BREAK", "CALL", "RET", "EXIT", "BAR", "WARPSYNC", "NANOSLEEP", "YIELD", "BRX", "NOP"})]
inside = False; fn = None; hist = collections.OrderedDict()
for line in open(path, errors="replace"):
    if line.startswith(".text."):
        inside = kern in line; fn = "kernel body (everything inlined)"; continue
    if line.startswith("//---") and ".text." in line: inside = False
    if not inside: continue
    m = re.match(r"^\$\S+\$_ZN3rtd\d+([A-Za-z_]+)", line)
    if m: fn = m.group(1) + " (out of line)"; continue
    m = re.match(r"^\$\S+\$(__internal_\d+_\$)?(\S+?):", line)
    if m: fn = "CUDA math slow paths (division / sqrt in FP64)"; continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P[0-9T]+\s+)?([A-Z0-9_]+)", line)
    if m: hist.setdefault(fn, collections.Counter())[m.group(1)] += 1
for fn, c in hist.items():
    tot = sum(c.values()); print(f"== {fn}: {tot} SASS instructions")
    rest = collections.Counter(c)
    for name, ops in GROUPS:
        n = sum(c[o] for o in ops); [rest.pop(o, None) for o in ops]
        detail = ", ".join(f"{o} {c[o]}" for o in sorted(ops, key=lambda o: -c[o]) if c[o])[:150]
        print(f"   {name:30s} {n:5d} {100 * n / tot:5.1f}%   {detail}")
    if rest: print(f"   {'other':30s} {sum(rest.values()):5d} {100 * sum(rest.values()) / tot:5.1f}%   " + ", ".join(f"{o} {n}" for o, n in rest.most_common(8)))
