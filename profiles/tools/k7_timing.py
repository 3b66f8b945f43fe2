#!/usr/bin/env python
"""Where the single-pod kernel (K7) spends its time: cycle stamps of its last CTA (profiling build, -DNGS_K7_TIMING).
usage: NGS_DEALER_LIB=profiles/tools/libngs_k7timing.so python profiles/tools/k7_timing.py"""
import ctypes as C
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import nano_gpu_scheduler_b200 as ngs  # noqa: E402
from nano_gpu_scheduler_b200 import workloads as wl  # noqa: E402

lib = ngs._abi.load()
N = 100000
gcount, free = wl.make_cluster(N, 0.5, 7)
tables = wl.make_load_tables(N, 7, 1_790_000_000)
load = ngs.make_load_ctx(wl.DEFAULT_PERIODS, 1_790_000_000 * 10**9)
d = ngs.GpuDealer(N, device=0)
d.snapshot_load(gcount, free)
for m, (value, ts, absent) in enumerate(tables):
    ok = np.repeat(~absent[:, None], 8, axis=1).astype(np.uint8)
    d.load_update_bulk(m, value, ts, value_ok=ok, ts_ok=ok)
pc, nc = wl.make_pods(400, 2, 4, 2, 10, 5, max_c=4)
pol = np.zeros(400, dtype=np.int32)
vp = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
o_node, o_score, o_gidx = np.zeros(1, dtype=np.int32), np.zeros(1, dtype=np.int32), np.zeros(4, dtype=np.int32)
st = np.zeros(5, dtype=np.int32)
rows, wall = [], []
for lref, label in ((C.c_void_p(C.addressof(load)), "load on"), (None, "load off")):
    rows, wall = [], []
    for s in range(400):
        t0 = time.perf_counter()
        rc = lib.ngs_schedule_batch(d._h, C.c_uint32(1), C.c_void_p(pc.ctypes.data + s * 16), C.c_void_p(nc.ctypes.data + s * 4), C.c_int(4),
                                    C.c_void_p(pol.ctypes.data + s * 4), lref, C.c_uint64(1), C.c_uint64(s), vp(o_node), vp(o_score), vp(o_gidx))
        wall.append(time.perf_counter() - t0)
        assert rc == 0
        if hasattr(lib, "ngs_debug_k7"):
            lib.ngs_debug_k7(d._h, vp(st))
            rows.append(st.copy())
    print(label, "wall p50 us", round(float(np.percentile(np.array(wall[50:]) * 1e6, 50)), 2))
    if rows:
        r = np.median(np.array(rows[50:]), axis=0) / 1.965e3
        print("  last CTA, us from its entry: after prologue %.2f, after row loop %.2f, after ticket %.2f, after plan+stores %.2f" % (r[1], r[2], r[3], r[4]))
d.close()
