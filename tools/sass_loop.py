#!/usr/bin/env python
"""Opcode histogram of the hot loops of one kernel: python tools/sass_loop.py lib.so 'jacobi_fast_kernelIfLi4ELi6E' [--list]"""
import collections, re, subprocess, sys
lib, pat = sys.argv[1], sys.argv[2]
txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout.split("\n")
starts = [i for i, l in enumerate(txt) if "Function :" in l]
s0 = [i for i in starts if pat in txt[i]][0]
s1 = min([i for i in starts if i > s0] + [len(txt)])
ins = []
for l in txt[s0:s1]:
    m = re.match(r"\s*/\*([0-9a-f]{4,5})\*/\s+(.*?);", l)
    if m:
        ins.append((int(m.group(1), 16), m.group(2).strip()))
loops = []
for a, t in ins:
    m = re.search(r"BRA(?:\.U)? (?:!?U?P\d, )?0x([0-9a-f]+)", t)
    if m and int(m.group(1), 16) < a:
        loops.append((int(m.group(1), 16), a))
print("instructions", len(ins), "loops", [(hex(a), hex(b), (b - a) // 16 + 1) for a, b in loops])
def op(t):
    w = t.split()
    return (w[1] if w[0].startswith("@") else w[0]).split(".")[0]
for a0, a1 in loops:
    body = [(a, t) for a, t in ins if a0 <= a <= a1]
    if len(body) < 100:
        continue
    c = collections.Counter(op(t) for _, t in body)
    print(hex(a0), len(body), c.most_common())
    if "--list" in sys.argv:
        for a, t in body:
            if op(t) not in ("FADD", "FADD2", "FMUL", "SHFL", "DADD", "DMUL"):
                print("   ", hex(a), t)
