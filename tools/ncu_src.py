#!/usr/bin/env python
"""Summarise `ncu --page source --csv` output: instructions executed and stall samples by loop trip count
and opcode.   ncu -i X.ncu-rep --page source --csv > src.csv; python tools/ncu_src.py src.csv [min_exec]"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hdr_i]
ix = {n: i for i, n in enumerate(hdr)}
data = [r for r in rows[hdr_i + 1:] if len(r) == len(hdr)]
def num(r, n):
    try: return float(r[ix[n]])
    except Exception: return 0.0
tot_exec = sum(num(r, "Instructions Executed") for r in data)
tot_samp = sum(num(r, "# Samples") for r in data)
print(f"{len(data)} SASS instructions, {tot_exec:.0f} warp-instructions executed, {tot_samp:.0f} samples")
# group by execution count (a loop body shows up as one count)
by = collections.defaultdict(lambda: [0, 0.0, 0.0, collections.Counter()])
for r in data:
    e = num(r, "Instructions Executed")
    g = by[e]; g[0] += 1; g[1] += e; g[2] += num(r, "# Samples"); g[3][r[ix["Source"]].split()[0 if not r[ix["Source"]].startswith("@") else 1].split(".")[0]] += 1
print("exec-count  n_instr  share_exec  share_samples  top opcodes")
for e, g in sorted(by.items(), key=lambda kv: -kv[1][1])[:14]:
    print(f"{e:10.0f} {g[0]:8d} {100*g[1]/tot_exec:9.1f}% {100*g[2]/max(tot_samp,1):12.1f}%  {dict(g[3].most_common(8))}")
stalls = [n for n in hdr if n.startswith("stall_") and "Not Issued" not in n]
tot = {n: sum(num(r, n) for r in data) for n in stalls}
print("stalls:", {k: int(v) for k, v in sorted(tot.items(), key=lambda kv: -kv[1])[:8]})
if len(sys.argv) > 2:
    lo = float(sys.argv[2])
    print("--- hottest instructions by samples")
    for r in sorted(data, key=lambda r: -num(r, "# Samples"))[:40]:
        if num(r, "Instructions Executed") >= lo:
            top = sorted(((num(r, n), n) for n in stalls), reverse=True)[:2]
            print(r[ix["Address"]][-5:], f'{num(r,"# Samples"):6.0f} {num(r,"Instructions Executed"):8.0f}', r[ix["Source"]][:90], [(n, int(v)) for v, n in top if v])
