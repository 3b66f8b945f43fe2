import ctypes as C, time, sys, os
import numpy as np
sys.path.insert(0, os.getcwd())
import nano_gpu_scheduler_b200 as ngs
from nano_gpu_scheduler_b200 import workloads as wl
import bench
lib = ngs._abi.load()
cfg = bench.CONFIGS["C3"] if hasattr(bench, "CONFIGS") else None
N = 100000
gcount, free = wl.make_cluster(N, 0.5, 7)
tables = wl.make_load_tables(N, 7, 1_790_000_000)
load = ngs.make_load_ctx(wl.DEFAULT_PERIODS, 1_790_000_000 * 10**9)
vp = lambda a: None if a is None else a.ctypes.data_as(C.c_void_p)
for n in (1000, 100000):
    d = ngs.GpuDealer(n, device=0)
    d.snapshot_load(gcount[:n], free[:n])
    for m, (value, ts, absent) in enumerate(tables):
        ok = np.repeat(~absent[:n, None], 8, axis=1).astype(np.uint8)
        d.load_update_bulk(m, value[:n], ts[:n], value_ok=ok, ts_ok=ok)
    feas = np.zeros(n, dtype=np.uint8); reason = np.zeros(n, dtype=np.int8); score = np.zeros(n, dtype=np.int32)
    dem = np.array([30, 20, 10], dtype=np.int32)
    lref = C.c_void_p(C.addressof(load))
    for label, args in (("feas+reason", (feas, reason, None)), ("feas", (feas, None, None)), ("none", (None, None, None)), ("feas+reason+score", (feas, reason, score))):
        ts_ = []
        for it in range(120):
            dem[0] = 5 * (it % 10 + 1)
            t0 = time.perf_counter()
            rc = lib.ngs_assume(d._h, None, C.c_uint32(n), vp(dem), C.c_int(3), C.c_int(0), lref, vp(args[0]), vp(args[1]), vp(args[2]), None, None)
            ts_.append(time.perf_counter() - t0)
            assert rc == 0
        print(n, "load-on", label, "p50 us", round(float(np.percentile(np.array(ts_[20:]) * 1e6, 50)), 1))
    for label, args in (("feas+reason", (feas, reason, None)), ("none", (None, None, None))):
        ts_ = []
        for it in range(120):
            dem[0] = 5 * (it % 10 + 1)
            t0 = time.perf_counter()
            rc = lib.ngs_assume(d._h, None, C.c_uint32(n), vp(dem), C.c_int(3), C.c_int(0), None, vp(args[0]), vp(args[1]), vp(args[2]), None, None)
            ts_.append(time.perf_counter() - t0)
        print(n, "load-off", label, "p50 us", round(float(np.percentile(np.array(ts_[20:]) * 1e6, 50)), 1))
    ts_ = []
    for it in range(120):
        t0 = time.perf_counter()
        lib.ngs_score(d._h, None, C.c_uint32(n), vp(dem), C.c_int(3), C.c_int(0), None, vp(score))
        ts_.append(time.perf_counter() - t0)
    print(n, "score (cached)", round(float(np.percentile(np.array(ts_[20:]) * 1e6, 50)), 1))
    a = np.zeros(n, dtype=np.int32); b = np.ones(n, dtype=np.int32)
    t0 = time.perf_counter()
    for _ in range(100): np.copyto(a, b)
    print(n, "numpy copy of 4n bytes us", round((time.perf_counter() - t0) * 1e4, 1))
    d.close()
