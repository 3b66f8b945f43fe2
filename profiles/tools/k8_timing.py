#!/usr/bin/env python
"""Where a request to the resident kernel (K8) spends its time: %globaltimer stamps (profiling build, -DNGS_K8_TIMING),
nanoseconds since CTA 0 saw the request.  usage: NGS_DEALER_LIB=profiles/tools/libngs_k8timing.so python profiles/tools/k8_timing.py"""
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
o1, o2, o3 = C.c_int32(0), C.c_int32(0), np.zeros(4, dtype=np.int32)
st = np.zeros(5, dtype=np.int32)
for lref, label in ((C.c_void_p(C.addressof(load)), "load on"), (None, "load off")):
    rows, wall = [], []
    for s in range(400):
        t0 = time.perf_counter()
        rc = lib.ngs_serve_pod(d._h, C.c_void_p(pc.ctypes.data + s * 16), C.c_int(int(nc[s])), C.c_int(4), C.c_int(0), lref, C.c_uint64(1), C.c_uint64(s),
                               C.byref(o1), C.byref(o2), o3.ctypes.data_as(C.c_void_p))
        wall.append(time.perf_counter() - t0)
        assert rc == 0
        if hasattr(lib, "ngs_debug_k8"):
            lib.ngs_debug_k8(d._h, st.ctypes.data_as(C.c_void_p))
            rows.append(st.copy())
    print(label, "wall p50 us", round(float(np.percentile(np.array(wall[50:]) * 1e6, 50)), 2))
    if rows:
        r = np.median(np.array(rows[50:]), axis=0) / 1e3
        print("  answering CTA, us since CTA 0 saw the request: saw it published %.2f, request read + demand sorted %.2f, rows done %.2f, "
              "ticket %.2f, answer ready %.2f" % tuple(r))
d.close()
