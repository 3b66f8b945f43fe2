#!/usr/bin/env python
"""Per-phase cycle breakdown of k_schedule (profiling build with -DNGS_PHASE_TIMING).

usage (on a GPU box):  python profiles/tools/phase_timing.py [workload] [pods]
Builds nano-gpu-scheduler_b200/libngs_dealer_dbg.so, runs one batch and prints, for thread 0
(warp 0: control path) and thread 32 (a plain worker) the average cycles per pod per phase,
as min / median / max over CTAs."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
PKG = os.path.join(ROOT, "nano-gpu-scheduler_b200")
DBG = os.path.join(PKG, "libngs_dealer_dbg.so")
subprocess.run(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-shared", "-Xcompiler", "-fPIC",
                "-diag-suppress", "177", "-DNGS_PHASE_TIMING", "-o", DBG, os.path.join(PKG, "csrc", "ngs_dealer.cu"),
                os.path.join(PKG, "csrc", "ngs_parse.cpp")], check=True)
os.environ["NGS_DEALER_LIB"] = DBG
import bench  # noqa: E402
import nano_gpu_scheduler_b200 as ngs  # noqa: E402
from nano_gpu_scheduler_b200 import workloads as wl  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "C3"
P = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
cfg = bench.workload_spec(name)
gcount, free, pc, nc, tables = bench.build_inputs(cfg, 0, 1, 2 * P)
d = ngs.GpuDealer(cfg["nodes"])
d.snapshot_load(gcount, free)
load = None
if cfg["load"]:
    for m, (value, ts, absent) in enumerate(tables):
        ok = np.repeat(~absent[:, None], 8, axis=1).astype(np.uint8)
        d.load_update_bulk(m, value, ts, value_ok=ok, ts_ok=ok)
    load = ngs.make_load_ctx(wl.DEFAULT_PERIODS, bench.NOW_NS)
pol = np.full(2 * P, cfg["policy"], dtype=np.int32)
d.schedule_batch(pc[:P], nc[:P], pol[:P], load=load)
d.schedule_batch(pc[P:], nc[P:], pol[P:], load=load)
lib = ngs._abi.load()
lib.ngs_debug_phases.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(C.c_uint32)]
buf = np.zeros(4096 + 256 * 256 * 8, dtype=np.uint64)
grid = C.c_uint32(0)
assert lib.ngs_debug_phases(d._h, buf.ctypes.data_as(C.c_void_p), C.byref(grid)) == 0
g = grid.value
R = P / 2.0  # rounds: two arg-max pods per grid exchange (bench workloads hold no random-policy pods)
a = buf[: 16 * g].reshape(g, 16).astype(np.float64) / R
names = ["-", "dirty-row fix-ups (bits, one warp's keys recomputed)", "combine + exact plan (cw0: pod a + F; cw1: first atomicMax issued, plan of pod b's row)",
         "post (cw0: atomicMax xa + counter; cw1: second atomicMax + counter)", "cw1: poll the round record + resolve", "pair barrier (cw0 waits for the resolution)",
         "bind + the changed row's verdicts for the next round", "CTA barrier at the end of the previous round -> start of this one"]
print(f"{name}: {cfg['nodes']} nodes, {P} pods, grid {g}; cycles per ROUND of two pods (min / median / max over CTAs)")
for who, off in (("control warp 0 (pod a: plan, post, bind)", 0), ("control warp 1 (pod b: top two, poll, bind)", 8)):
    print(who)
    for i in (7, 1, 2, 3, 4, 5, 6):
        col = a[:, off + i]
        print(f"  {names[i]:100s} {col.min():8.0f} {np.median(col):8.0f} {col.max():8.0f}")
    print(f"  {'total':14s} {a[:, off:off + 8].sum(axis=1).mean():8.0f}")


w = buf[4096: 4096 + 32 * g].reshape(g, 32).astype(np.float64) / R
nw = int((w[0] > 0).sum())
per_cta = w[:, :nw].max(axis=1)
print(f"workers: filter of the NEXT round's pods, cycles per round, slowest warp of each CTA ({nw} worker warps): min {per_cta.min():.0f} median {np.median(per_cta):.0f} "
      f"max {per_cta.max():.0f} (CTA {int(per_cta.argmax())}); within the slowest CTA: fastest warp {w[int(per_cta.argmax()), :nw].min():.0f}")
order = np.argsort(-per_cta)[:6]
print("slowest CTAs:", [(int(c), int(per_cta[c])) for c in order])
e = buf[20000:20016].astype(np.float64)
rounds = e[0] / g if e[0] else 1.0  # (event 0 is counted by every CTA)
print(f"events over {rounds:.0f} two-pod rounds: announcement needed {e[1] / g:.0f}; F of the winner: no {e[8] / g:.0f}, yes {e[9] / g:.0f}, open {e[10] / g:.0f}, no winner {e[11] / g:.0f}")
print(f"  owner of pod a: {e[3]:.0f} binds, {e[4] / max(e[3], 1):.0f} cycles from resolution to the end of its bind phase; pod b on pod a's row: {e[2]:.0f}")
print(f"  owner of pod b (other row): {e[5]:.0f} binds, {e[6] / max(e[5], 1):.0f} cycles; of them planned after the fact (second-best row): {e[7]:.0f}")
# ---- per-round trace (GPU-wide nanoseconds): who posts last, how long the exchange takes once everybody has posted ----
nr = min(256, int(R))
tr = buf[32768: 32768 + nr * g * 8].reshape(nr, g, 8).astype(np.float64)
tr = tr[8:]  # skip the first rounds
post_a, plan_done, bind0, start1, post_b, resolved, bind1 = (tr[:, :, k] for k in (0, 1, 2, 3, 4, 5, 6))
last_post = np.maximum(post_a, post_b).max(axis=1)
first_post = np.minimum(post_a, post_b).min(axis=1)
res_med = np.median(resolved, axis=1)
print(f"trace over {len(tr)} rounds (ns, 1 ns = 1.965 cycles): round period {np.median(np.diff(res_med)):.0f}")
print(f"  spread of the posts (last - first CTA) median {np.median(last_post - first_post):.0f}; resolution after the LAST post median {np.median(res_med - last_post):.0f}")
print(f"  median CTA: round start -> its posts {np.median(np.median(np.maximum(post_a, post_b) - start1, axis=1)):.0f}; last CTA: {np.median((np.maximum(post_a, post_b) - start1).max(axis=1)):.0f}")
print(f"  previous resolution -> round start (binds, CTA barrier): median CTA {np.median(np.median(start1[1:] - resolved[:-1], axis=1)):.0f}, slowest CTA {np.median((start1[1:] - resolved[:-1]).max(axis=1)):.0f}")
print(f"  plan of pod a (post -> plan done) median CTA {np.median(plan_done - post_a):.0f}, p95 {np.percentile(plan_done - post_a, 95):.0f}; plan done after the resolution in {100 * np.mean(plan_done > resolved):.0f} % of (round, CTA)")
late = np.argmax(np.maximum(post_a, post_b), axis=1)
prev_owner_hint = (bind0[:-1] - resolved[:-1]).argmax(axis=1)
print(f"  the last poster is the CTA with the longest previous bind phase in {100 * np.mean(late[1:] == prev_owner_hint):.0f} % of rounds")
