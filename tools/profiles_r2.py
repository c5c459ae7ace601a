#!/usr/bin/env python
"""Turn what tools/capture_profiles.sh r2 left in gpurun_out/ into the tracked files under profiles/:
bench_r2.json, bench_reference_arm_r2.json, launches_r2.csv, ncu_traffic_r2.json and README_r2.md (sections 1-4
from the captures; sections 5-6 from profiles/experiments_r2.md and profiles/scaling_r2.jsonl).
    python tools/profiles_r2.py"""
import collections
import csv
import json
import os
import re
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

G = os.path.join(ROOT, "gpurun_out")
P = os.path.join(ROOT, "profiles")


def short(name):
    return re.sub(r"\(.*", "", name).replace("void <unnamed>::", "").strip()


def raw_page(x):
    rows = list(csv.reader(open(os.path.join(G, f"prof_step_{x}_r2.raw.csv"))))
    hdr, units = rows[0], rows[1]
    return {short(r[hdr.index("Kernel Name")]): {h: (r[i], units[i]) for i, h in enumerate(hdr)} for r in rows[2:]}


def traffic_table(K):
    tot = {}
    for name, m in K.items():
        mm = re.search(r"jacobi_\w*kernel<float, 4, (\d)", name)
        if mm:
            b = (float(m["dram__bytes_read.sum"][0]) + float(m["dram__bytes_write.sum"][0])) * 1e9
            tot[int(mm.group(1))] = tot.get(int(mm.group(1)), 0) + b
    table = {"_comment": "dram__bytes_read.sum + dram__bytes_write.sum per Jacobi launch at 16384x16384 (fast + edge + wall kernel of the "
                         "split launch summed), from the ncu --set full captures of tools/capture_profiles.sh r2; bench.py averages the "
                         "depths a step uses into roofline.traffic and drops the table when source_sha1 no longer matches the kernel sources",
             "source_sha1": bench.kernel_source_hash()}
    for k, b in sorted(tot.items(), reverse=True):
        table[f"jacobi_fast_kernel<float,4,{k}>@16384x16384"] = round(b)
    json.dump(table, open(os.path.join(P, "ncu_traffic_r2.json"), "w"), indent=1)
    return tot


def launch_table(d):
    rows = [r for r in csv.reader(open(os.path.join(P, "launches_r2.csv"))) if len(r) > 5]
    h = [i for i, r in enumerate(rows) if r[0] == "ID"][0]
    H, data = rows[h], rows[h + 1:]
    ki, vi, ui = H.index("Kernel Name"), H.index("Metric Value"), H.index("Metric Unit")
    agg = collections.OrderedDict()
    for r in data:
        if int(r[0]) >= 206:      # init + five timesteps of 41 kernels; the rest is the end-to-end leg
            continue
        v = float(r[vi].replace(",", "")) * {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(r[ui], 1.0)
        a = agg.setdefault(short(r[ki]), [0, 0.0])
        a[0] += 1
        a[1] += v
    t = sum(a[1] for k, a in agg.items() if "init" not in k)
    L = ["| kernel | launches | total ms | mean us | share of the five steps |", "|---|---|---|---|---|"]
    jac = 0.0
    for k, (n, v) in agg.items():
        if "init" in k:
            continue
        L.append(f"| `{k}` | {n} | {v / 1e3:.3f} | {v / n:.1f} | {v / t:.1%} |")
        jac += v if "jacobi" in k else 0.0
    L.append(f"\nJacobi kernels: {jac / t:.1%} of the captured time (serialised: the edge and wall kernels do not overlap the fast kernel under "
             f"ncu); {d['roofline']['share_of_step']:.1%} of the event-timed step in `bench_r2.json`: the shares agree.\n")
    return L


WANT = [("gpu__time_duration.sum", "duration"), ("dram__bytes_read.sum", "DRAM read"), ("dram__bytes_write.sum", "DRAM written"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM % of ncu peak"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue slots busy %"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps active %"), ("launch__registers_per_thread", "registers"),
        ("launch__grid_size", "CTAs"), ("smsp__inst_executed.sum", "warp instructions"),
        ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "stall long_scoreboard"),
        ("smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio", "stall math_pipe_throttle"),
        ("smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "stall wait"),
        ("smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio", "stall not_selected"),
        ("smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "stall short_scoreboard")]
NAMES = ["predictor_kernel<float, 4, 2>", "jacobi_fast_kernel<float, 4, 8>", "jacobi_fast_kernel<float, 4, 7>",
         "jacobi_edge_kernel<float, 4, 8>", "jacobi_kernel<float, 4, 8, 0>", "corrector_kernel<float, 4, 0, 2>"]


def ncu_table(K):
    L = ["| metric | " + " | ".join(f"`{n}`" for n in NAMES) + " |", "|---|" + "---|" * len(NAMES)]
    for m, label in WANT:
        cells = []
        for n in NAMES:
            v, u = K.get(n, {}).get(m, ("", ""))
            try:
                v = f"{float(v):.4g}"
            except ValueError:
                pass
            cells.append(f"{v} {u}".strip())
        L.append(f"| {label} | " + " | ".join(cells) + " |")
    return L


def scaling_table():
    L = ["| N | weak: Gcell-updates/s (ms per step) | strong 32768^2: ms per step | one GPU, same run (ms) | speed-up | host time to issue a "
         "step (ms) | e2e ms per step (host aggregate GB/s) |", "|---|---|---|---|---|---|---|"]
    for line in open(os.path.join(P, "scaling_r2.jsonl")):
        d = json.loads(line)
        s = d["config"]["strong"]
        L.append(f"| {d['n_gpus']} | {d['value']:.1f} ({d['ms_per_step']:.3f}) | {s['ms_per_step']:.3f} | {s['one_gpu_ms_per_step']:.2f} | "
                 f"{s['speedup_vs_one_gpu']:.2f} x ({s['efficiency']:.1%}) | {d['config']['host_issue_ms_per_step']:.2f} | "
                 f"{d['e2e']['ms_per_step']:.0f} ({d['e2e']['host_link_GBps_aggregate']:.0f}) |")
    return L


def main():
    for f in ("bench_r2.json", "bench_reference_arm_r2.json", "launches_r2.csv"):
        shutil.copy(os.path.join(G, f), os.path.join(P, f))
    K = {**raw_page("a"), **raw_page("b")}
    tot = traffic_table(K)
    d = json.load(open(os.path.join(P, "bench_r2.json")))
    ref = json.load(open(os.path.join(P, "bench_reference_arm_r2.json")))
    r, ok, e, cb, g = d["roofline"], d["roofline"]["other_kernels"], d["e2e"], d["cpu_baseline"], d["headline_general_spacing"]
    comp = r["compulsory_bytes_per_launch"]
    L = [open(os.path.join(P, "parts_r2", "head.md")).read()]
    L.append("## 1. Bench line (bench_r2.json)\n")
    L.append(f"* workload `{d['config']['workload']}`; 13 Jacobi launches per step (9 of depth 8, 4 of depth 7); arithmetic mode 2 (power-of-two spacing)")
    L.append(f"* **value = {d['value']:.2f} Gcell-updates/s** ({d['ms_per_step']:.3f} ms per timestep; {d['gpu_launches'] // d['steps']} kernels per step); clocks {d['clocks']}")
    L.append(f"* Jacobi: {r['launch_ms'] * 1e3:.0f} us per mean launch = {r['gcell_sweeps_per_s']:.0f} Gcell-sweeps/s, {r['share_of_step']:.1%} of the step. "
             f"`roofline.achieved` = compulsory 12 B/cell per launch / launch time = {r['achieved']:.0f} GB/s = **{r['frac']:.3f} of the measured copy peak** "
             f"({r['peak']} GB/s); algorithmic (12 B/cell per sweep) {r['algorithmic']['achieved']:.0f} GB/s = {r['algorithmic']['frac']:.2f} x the peak; "
             f"ncu bytes per launch: {tot[8] / 1e9:.3f} GB (depth 8), {tot[7] / 1e9:.3f} GB (depth 7) = {tot[8] / comp:.2f} / {tot[7] / comp:.2f} x the compulsory "
             f"{comp / 1e9:.3f} GB; `roofline.traffic` in a bench run made with these sources: {r['traffic']}")
    L.append(f"* predictor {ok['predictor_ms']:.3f} ms ({ok['predictor_GBps']:.0f} GB/s = {ok['predictor_frac']:.3f}), corrector {ok['corrector_ms']:.3f} ms "
             f"({ok['corrector_GBps']:.0f} GB/s = {ok['corrector_frac']:.3f}); whole step: {r['whole_step']['GBps']:.0f} GB/s of compulsory traffic = {r['whole_step']['frac']:.3f}")
    L.append(f"* the same step with a spacing that is not a power of two (`FLUIDB200_DIVMODE=1`): {g['ms_per_step']:.2f} ms = {g['value']:.1f} Gcell-updates/s")
    L.append(f"* e2e (`fluidb200_migrate_solve`, pinned host arrays, 3.2 GB each way per step): **{e['value']:.2f} Gcell-updates/s** ({e['ms_per_step']:.1f} ms/step, "
             f"{e['link_GBps_each_way']:.1f} GB/s in each direction at once); as separate upload / solve / download calls (round 1's path): "
             f"{e['separate_calls']['ms_per_step']:.1f} ms; two solver handles on two host threads: {e['two_instances']['ms_per_step']:.1f} ms per step")
    L.append(f"* cpu_baseline: reference threaded C++ (8 threads) {cb['value'] * 1e3:.2f} Mcell-updates/s on a 1536^2 sample; serial (1 core) "
             f"{cb['serial']['value'] * 1e3:.2f} Mcell-updates/s on a 768^2 sample; reference arm (1024^2 sample): threaded {ref['value'] * 1e3:.2f}, "
             f"serial {ref['cpu_baseline']['serial']['value'] * 1e3:.2f} Mcell-updates/s")
    L.append("\n## 2. BASELINE configs[0] and [1] (1000 timesteps x 50 sweeps), ms for the whole run\n")
    L.append("| grid | resident solver (default) | streaming + CUDA graphs | streaming, plain launches | reference serial C++ (1 core) | reference threaded C++ "
             "(8 threads) | reference CUDA, sm_100a build, same GPU |\n|---|---|---|---|---|---|---|")
    for c in (d["config0_41x41"], d["config1_1000x1000"]):
        ex = lambda k: f"{c[k]['ms_total']:.0f}" + (" (extrapolated from %d steps)" % c[k]["steps_run"] if c[k].get("extrapolated") else "")
        L.append(f"| {c['grid']} | **{c['resident']['ms_total']:.1f}** ({c['resident']['kernel_launches']} launch) | {c['streaming_graph']['ms_total']:.0f} "
                 f"({c['streaming_graph']['kernel_launches']} launches) | {c['streaming_plain']['ms_total']:.0f} | {ex('reference_cpu_serial')} | "
                 f"{ex('reference_cpu_threaded')} | {ex('reference_cuda_sm100a')} |")
    L.append(f"\n41^2: v bit-equal to the reference's serial solver run in the same process: {d['config0_41x41'].get('v_bit_equal_to_reference_serial')}. "
             "1000^2 uses a stable time step (the reference's 0.001 diverges there).")
    L.append(open(os.path.join(P, "parts_r2", "resident.md")).read())
    L.append("## 3. Launch list (launches_r2.csv), first 206 kernels: share of the step\n")
    L += launch_table(d)
    L.append("## 4. `ncu --set full`, one launch of each kernel of a timestep (16384^2)\n")
    L += ncu_table(K)
    L.append(open(os.path.join(P, "parts_r2", "reading.md")).read())
    L.append(open(os.path.join(P, "parts_r2", "experiments.md")).read())
    L.append("## 6. Scaling (`bench.py --gpus N`, one process per GPU; scaling_r2.jsonl)\n")
    L += scaling_table()
    L.append(open(os.path.join(P, "parts_r2", "scaling.md")).read())
    open(os.path.join(P, "README_r2.md"), "w").write("\n".join(L))
    print("wrote profiles/README_r2.md, ncu_traffic_r2.json", {k: round(v / 1e9, 3) for k, v in tot.items()})


if __name__ == "__main__":
    main()
