#!/usr/bin/env python
"""Turn the scratch files of a gpurun call (gpurun_out/) into the committed summaries under
profiles/<round>/.   usage: python profiles/tools/summarize.py r01"""
import collections
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rnd = sys.argv[1] if len(sys.argv) > 1 else "r01"
G = os.path.join(ROOT, "gpurun_out")
OUT = os.path.join(ROOT, "profiles", rnd)
os.makedirs(OUT, exist_ok=True)
MULT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def last_json(path):
    return json.loads([l for l in open(path) if l.startswith("{")][-1])


def raw_metrics(rep, want):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rr = list(csv.reader(io.StringIO(raw)))
    hdr, units, vals = rr[0], rr[1], rr[2]
    return {w: (vals[i], units[i]) for w in want for i, h in enumerate(hdr) if h == w}


WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic",
        "smsp__inst_executed.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct"]


def dram_bytes(m):
    f = lambda k: float(m[k][0].replace(",", "")) * MULT[m[k][1]]  # noqa: E731
    return f("dram__bytes_read.sum") + f("dram__bytes_write.sum")


b = last_json(os.path.join(G, f"bench_{rnd}_n1.log"))
json.dump(b, open(os.path.join(OUT, "bench_n1.json"), "w"), indent=1)
rows = [r for r in csv.DictReader(l for l in open(os.path.join(G, f"launches_{rnd}.csv")) if not l.startswith("=="))]
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows:
    if r.get("Metric Name") == "gpu__time_duration.sum":
        k = r["Kernel Name"][:70]
        agg[k][0] += 1
        agg[k][1] += float(r["Metric Value"].replace(",", ""))
tot = sum(v[1] for v in agg.values())
with open(os.path.join(OUT, "launch_list_summary.txt"), "w") as f:
    f.write("ncu --metrics gpu__time_duration.sum --clock-control none -c 200 : python bench.py --steps 2 --warmup 3 --pods-per-step 512 "
            "--no-cpu-baseline --skip-e2e --skip-extras\n(cold-cache, serialised launches: compare shares, not absolutes)\n"
            "(a step = k_prep_demands + k_schedule; the k_schedule_one launches are the single-pod latency arm that runs after the timed steps,\n"
            " the fill kernels are the L2 flush between steps)\n")
    for k, v in sorted(agg.items(), key=lambda x: -x[1][1]):
        f.write(f"{v[1] / 1e3:12.1f} us  {100 * v[1] / tot:5.1f}%  n={v[0]:3d}  {k}\n")
m = raw_metrics(os.path.join(G, f"prof_sched_{rnd}.ncu-rep"), WANT)
json.dump(m, open(os.path.join(OUT, "k_schedule_ncu_full_summary.json"), "w"), indent=1)
json.dump({"kernel": "k_schedule", "dram_bytes_per_launch": dram_bytes(m), "pods_per_launch": 512, "nodes": 100000,
           "algorithmic_bytes_per_launch": 512 * 100000 * 125,
           "source": f"profiles/{rnd}/k_schedule_ncu_full_summary.json (ncu --set full --clock-control none, 512 pods per launch)"},
          open(os.path.join(ROOT, "profiles", "k_schedule_traffic.json"), "w"), indent=1)
k1 = os.path.join(G, f"prof_k1_on_{rnd}.ncu-rep")
if os.path.exists(k1):
    m1 = raw_metrics(k1, WANT)
    m1["dram_bytes_per_eval"] = (dram_bytes(m1) / 4_000_000, "byte")
    json.dump(m1, open(os.path.join(OUT, "k_assume_load_on_ncu_full_summary.json"), "w"), indent=1)
k0 = os.path.join(G, f"prof_k1_off_{rnd}.ncu-rep")
if os.path.exists(k0):
    m0 = raw_metrics(k0, WANT)
    m0["dram_bytes_per_eval"] = (dram_bytes(m0) / 4_000_000, "byte")
    json.dump(m0, open(os.path.join(OUT, "k_scan_tma_load_off_ncu_full_summary.json"), "w"), indent=1)
kb = os.path.join(G, f"k1_bench_{rnd}.log")
if os.path.exists(kb):
    json.dump([json.loads(l) for l in open(kb) if l.startswith("{")], open(os.path.join(OUT, "k_assume_roofline.json"), "w"), indent=1)
hot = subprocess.run([sys.executable, os.path.join(ROOT, "profiles", "tools", "ncu_lines.py"), os.path.join(G, f"prof_sched_{rnd}.ncu-rep"),
                      os.path.join(ROOT, "nano-gpu-scheduler_b200", "libngs_dealer.so"), "k_schedule<0, 0>", "25"], capture_output=True, text=True).stdout
open(os.path.join(OUT, "k_schedule_hot_lines.txt"), "w").write(hot)
print(open(os.path.join(OUT, "launch_list_summary.txt")).read())
print({k: b[k] for k in ("value", "ms_per_step", "gpu_launches", "clocks")}, b["e2e"], b["latency_us"], b["roofline"]["frac"], b.get("cpu_baseline", {}).get("value"))
print(m)
