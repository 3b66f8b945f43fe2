#!/usr/bin/env python
"""Roofline of the streaming scan kernel K1 (k_assume): one filter+prioritize pass over N nodes,
outputs left in device memory, timed with CUDA events on the handle's stream, L2 flushed before
every launch.  usage: python profiles/tools/k1_bench.py [nodes] [reps]"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import nano_gpu_scheduler_b200 as ngs  # noqa: E402
from nano_gpu_scheduler_b200 import workloads as wl  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 16_000_000
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 20
peak, src = bench.measured_peak()
gcount, free = wl.make_cluster(n, 0.5, 0xB2000004)
tables = wl.make_load_tables(n, 0xB2000004, bench.NOW_S)
stream = torch.cuda.Stream()
d = ngs.GpuDealer(n, device=0, stream=stream.cuda_stream)
d.snapshot_load(gcount, free)
for m, (value, ts, absent) in enumerate(tables):
    ok = np.repeat(~absent[:, None], 8, axis=1).astype(np.uint8)
    d.load_update_bulk(m, value, ts, value_ok=ok, ts_ok=ok)
load = ngs.make_load_ctx(wl.DEFAULT_PERIODS, bench.NOW_NS)
feas = torch.empty(n, dtype=torch.uint8, device="cuda")
score = torch.empty(n, dtype=torch.int32, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
out = []
with torch.cuda.stream(stream):
    for name, demand, policy, ld, bpe in (("load off, C=1, binpack", [30], 0, None, 37), ("load off, C=3, spread", [40, 30, 20], 1, None, 37),
                                          ("load on, C=3, binpack", [40, 30, 20], 0, load, 125)):
        for _ in range(3):
            d.assume_dev(demand, policy, feas.data_ptr(), score.data_ptr(), load=ld)
        ms = []
        for _ in range(reps):
            flush.fill_(1)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            d.assume_dev(demand, policy, feas.data_ptr(), score.data_ptr(), load=ld)
            e1.record(stream)
            torch.cuda.synchronize()
            ms.append(e0.elapsed_time(e1))
        t = float(np.median(ms)) / 1e3
        res = dict(case=name, nodes=n, us=t * 1e6, evals_per_s=n / t, bytes_per_eval=bpe, achieved_gbs=n * bpe / t / 1e9, peak_gbs=peak,
                   frac=n * bpe / t / 1e9 / peak)
        out.append(res)
        print(json.dumps(res))
d.close()
