"""Shared helpers of the GPU parity tests: build the same snapshot on the oracle and on the GPU."""
import numpy as np

import oracle as o
from nano_gpu_scheduler_b200 import workloads as wl

NOW_S = 1_790_000_000  # fixed "time.Now()" of the load-aware tests
NOW_NS = NOW_S * 10**9 + 123_456_789


def make_pair(ngs, n, p_used, seed, hetero=False, sort_rule=o.SORT_GO116):
    gcount, free = wl.make_cluster(n, p_used, seed, hetero=hetero)
    oc = o.OracleCluster(n, sort_rule)
    oc.load_snapshot(gcount, free)
    d = ngs.GpuDealer(n, device=0)
    d.snapshot_load(gcount, free)
    return oc, d, gcount, free


def fill_load_tables(oc, d, n, seed, stale_frac=0.05, absent_frac=0.02):
    tables = wl.make_load_tables(n, seed, NOW_S, stale_frac, absent_frac)
    for m, (value, ts, absent) in enumerate(tables):
        ok = np.repeat(~absent[:, None], 8, axis=1).astype(np.uint8)
        d.load_update_bulk(m, value, ts, value_ok=ok, ts_ok=ok)
        # an absent node never got Add*Usage: same outcome as cells that do not parse
        oc.load_update_bulk(m, value, ts, ok=ok)
    return tables


def load_ctx_pair(ngs, periods=None, now_ns=NOW_NS):
    periods = periods or wl.DEFAULT_PERIODS
    return ngs.make_load_ctx(periods, now_ns), o.make_load_ctx(periods, now_ns)


def assert_same_snapshot(oc, d):
    gc, fr, rl = d.snapshot_read()
    ogc, ofr, orl = oc.snapshot()
    assert np.array_equal(gc, ogc)
    assert np.array_equal(fr, ofr)
    assert np.array_equal(rl, orl)
