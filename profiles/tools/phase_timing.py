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
a = buf[: 16 * g].reshape(g, 16).astype(np.float64) / P
names = ["cw0: exact plan only", "barrier (pod p settled)", "cw0: post-bind stats / cw1: atomicMax+bump", "control: iteration start -> pair barrier #2", "cw0: plan_group total / cw1: exchange wait", "pair barrier #1", "publish", "control: combine (best two per warp -> key)"]
print(f"{name}: {cfg['nodes']} nodes, {P} pods, grid {g}; cycles per pod (min / median / max over CTAs)")
for who, off in (("control warp 0 (plan, bind)", 0), ("control warp 1 (exchange)", 8)):
    print(who)
    for i in range(8):
        col = a[:, off + i]
        print(f"  {names[i]:84s} {col.min():8.0f} {np.median(col):8.0f} {col.max():8.0f}")
    print(f"  {'total':14s} {a[:, off:off + 8].sum(axis=1).mean():8.0f}")


w = buf[4096: 4096 + 32 * g].reshape(g, 32).astype(np.float64) / P
nw = int((w[0] > 0).sum())
per_cta = w[:, :nw].max(axis=1)
print(f"workers: filter of the NEXT pod, cycles per pod, slowest warp of each CTA ({nw} worker warps): min {per_cta.min():.0f} median {np.median(per_cta):.0f} "
      f"max {per_cta.max():.0f} (CTA {int(per_cta.argmax())}); within the slowest CTA: fastest warp {w[int(per_cta.argmax()), :nw].min():.0f}")
order = np.argsort(-per_cta)[:6]
print("slowest CTAs:", [(int(c), int(per_cta[c])) for c in order])
