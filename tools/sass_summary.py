#!/usr/bin/env python
"""profiles/sass_r2.md from the built library (no GPU needed): whole-function opcode counts of the hot kernels and the
opcode histogram of each one's largest loop.   python tools/sass_summary.py > profiles/sass_r2.md"""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "fluid_simulation_b200", "libfluidb200.so")
txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout.split("\n")
starts = [i for i, l in enumerate(txt) if "Function :" in l]
KERNELS = [("jacobi_fast_kernelIfLi4ELi8E", "jacobi_fast_kernel<float,4,8>"), ("jacobi_fast_kernelIfLi4ELi7E", "jacobi_fast_kernel<float,4,7>"),
           ("jacobi_edge_kernelIfLi4ELi8E", "jacobi_edge_kernel<float,4,8>"), ("jacobi_kernelIfLi4ELi8ELb0E", "jacobi_kernel<float,4,8,false>"),
           ("predictor_kernelIfLi4ELi2E", "predictor_kernel<float,4,2>"), ("predictor_kernelIfLi4ELi1E", "predictor_kernel<float,4,1>"),
           ("predictor_kernelIdLi2ELi1E", "predictor_kernel<double,2,1>"),
           ("corrector_kernelIfLi4ELb0ELi2E", "corrector_kernel<float,4,false,2>"), ("resident_kernelIfLi4ELi8E", "resident_kernel<float,4,8>"),
           ("resident_cell_kernelIfLi2E", "resident_cell_kernel<float,2>"), ("jacobi_fast_kernelIdLi2ELi6E", "jacobi_fast_kernel<double,2,6>")]


def body(pat):
    s0 = [i for i in starts if pat in txt[i]]
    if not s0:
        return []
    s0 = s0[0]
    s1 = min([i for i in starts if i > s0] + [len(txt)])
    ins = []
    for l in txt[s0:s1]:
        m = re.match(r"\s*/\*([0-9a-f]{4,5})\*/\s+(.*?);", l)
        if m:
            ins.append((int(m.group(1), 16), m.group(2).strip()))
    return ins


def op(t):
    w = t.split()
    return (w[1] if w[0].startswith("@") else w[0]).split(".")[0]


print("# profiles/sass_r2.md -- what the sm_100a cubin of `fluid_simulation_b200/libfluidb200.so` contains (round 2, final build)\n")
print("Made here (no GPU needed) by `tools/sass_summary.py` from `cuobjdump -sass`: whole-function opcode counts and the largest loop of each kernel.")
print("`FFMA` in a Jacobi kernel would mean a contracted multiply-add (it would change result bits): there is none. `FADD2` = Blackwell's packed")
print("fp32 add; `LDGSTS` = `cp.async`; `UBLKCP` / `UTMALDG` = TMA (none in the product build: the bulk-copy variant of the lean Jacobi loop was built")
print("and measured, `README_r2.md` section 5, and lost); `LDL`/`STL` = spills; `CALL` = out-of-line second evaluation of flagged cells.\n")
cols = ["FFMA", "FADD2", "FADD", "FMUL", "DFMA", "DMUL", "F2F", "SHFL", "LDGSTS", "UBLKCP+UTMALDG", "LDL+STL", "CALL", "WARPSYNC"]
print("| kernel | SASS instructions | " + " | ".join(cols) + " |")
print("|---|---|" + "---|" * len(cols))
bodies = {}
for pat, name in KERNELS:
    ins = body(pat)
    if not ins:
        continue
    bodies[name] = ins
    c = collections.Counter(op(t) for _, t in ins)
    vals = [c[k] if "+" not in k else sum(c[x] for x in k.split("+")) for k in cols]
    print(f"| `{name}` | {len(ins)} | " + " | ".join(str(v) for v in vals) + " |")
print("\n(FFMA in the predictor / corrector / resident kernels: `fmaf(-4, c, s)`, the exact evaluation of the reference's `s - 4.*c`, and the inlined")
print("IEEE-division sequences of the last-resort DIV_EXACT evaluation; DFMA in the fp64 predictor: the remainder corrections of its quotients.)\n")
print("## Largest loops (static counts; cold paths inside the loop included)\n")
for name, ins in bodies.items():
    loops = []
    for a, t in ins:
        m = re.search(r"BRA(?:\.U)? (?:!?U?P\d, )?0x([0-9a-f]+)", t)
        if m and int(m.group(1), 16) < a:
            loops.append((int(m.group(1), 16), a))
    best = None
    for a0, a1 in loops:
        b = [(a, t) for a, t in ins if a0 <= a <= a1]
        math = sum(1 for _, t in b if op(t) in ("FADD", "FADD2", "FMUL", "DADD", "DMUL", "DFMA", "FFMA"))
        if math >= 32 and (best is None or len(b) < len(best[2])) and math >= 0.5 * max(1, sum(1 for _, t in ins if op(t) in ("FADD", "FADD2", "FMUL", "DADD", "DMUL", "DFMA", "FFMA")) ) * 0.2:
            best = (a0, a1, b)
    if not best:
        continue
    a0, a1, b = best
    c = collections.Counter(op(t) for _, t in b)
    print(f"* `{name}`, loop {hex(a0)}-{hex(a1)}, {len(b)} instructions: " + ", ".join(f"{k} {v}" for k, v in c.most_common(16)))
print("\n`jacobi_fast_kernel<float,4,8>`: 8 steps x 8 levels x 4 cells per lane = 256 cell-sweeps per loop trip against 4.5 instructions of pure arithmetic")
print("per cell-sweep (2 FADD + 2 x 1/2 FADD2 + FMUL + 1/2 SHFL). Round 1's 4-step body: 777 instructions per 128 cell-sweeps = 6.07, of them 90 MOV")
print("(shift-register rotation the 4-step unroll could not rename away). The loop has no WARPSYNC (the function's sit in the pipeline fill / drain code).")
print("`predictor_kernel<float,4,2>`: no spills and no IEEE-division code any more: the second evaluation of a flagged row (`CALL`) is the double Laplacian")
print("tail (`DIV_POW2_WIDE`), a handful of instructions; until this round's last change it was `__fdiv_rn` / `__ddiv_rn` sequences with 16 bytes of spills around the calls.")
