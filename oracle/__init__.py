"""ctypes binding of the CPU oracle (oracle/dealer_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs, never by the product package.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libdealer_oracle.so")

POLICY_BINPACK, POLICY_SPREAD, POLICY_RANDOM = 0, 1, 2
SORT_GO116, SORT_GO119 = 0, 1
REASON_OK, REASON_UNKNOWN_NODE, REASON_INFEASIBLE = 0, 1, 2
METRIC_CORE, METRIC_MEMORY = 0, 1
MAX_GPUS, MAX_CONTAINERS, MAX_PERIODS = 64, 64, 8
INT32_MIN = -(2**31)


class LoadCtx(C.Structure):
    _fields_ = [
        ("n_periods", C.c_int32),
        ("metric", C.c_int32 * MAX_PERIODS),
        ("window_ns", C.c_int64 * MAX_PERIODS),
        ("now_ns", C.c_int64),
    ]


def make_load_ctx(periods, now_ns):
    """periods: list of (metric, window_ns) already resolved as getActiveDuration does."""
    L = LoadCtx()
    L.n_periods = len(periods)
    for i, (m, w) in enumerate(periods):
        L.metric[i] = m
        L.window_ns[i] = w
    L.now_ns = now_ns
    return L


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "dealer_oracle.c")
    hdr = os.path.join(_HERE, "dealer_oracle.h")
    stale = (
        force
        or not os.path.exists(_SO)
        or os.path.getmtime(_SO) < max(os.path.getmtime(src), os.path.getmtime(hdr))
    )
    if stale:
        subprocess.run(["make", "-s", "-B", "-C", _HERE], check=True)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(build())
        vp, u32, i32, i64, u64 = C.c_void_p, C.c_uint32, C.c_int32, C.c_int64, C.c_uint64
        L.or_create.restype = vp
        L.or_create.argtypes = [u32, C.c_int]
        L.or_destroy.argtypes = [vp]
        L.or_set_buggy_restore.argtypes = [vp, C.c_int]
        L.or_node_upsert.argtypes = [vp, u32, C.c_int]
        L.or_node_set.argtypes = [vp, u32, C.c_int, vp, vp]
        L.or_node_get.argtypes = [vp, u32, vp, vp, vp]
        L.or_snapshot_load.argtypes = [vp, u32, u32, vp, vp]
        L.or_snapshot_read.argtypes = [vp, u32, u32, vp, vp, vp]
        L.or_load_update_bulk.argtypes = [vp, C.c_int, u32, u32, vp, vp, vp]
        L.or_node_apply.argtypes = [vp, u32, vp, vp, C.c_int, C.c_int]
        L.or_load_update.argtypes = [vp, u32, C.c_int, C.c_int, C.c_double, C.c_int, i64, C.c_int]
        L.or_load_add_node.argtypes = [vp, u32, C.c_int]
        L.or_assume.argtypes = [vp, vp, u32, vp, C.c_int, C.c_int, vp, vp, vp, vp, vp, vp]
        L.or_bind.argtypes = [vp, u32, vp, C.c_int, C.c_int, vp, vp, vp]
        L.or_schedule_batch.argtypes = [vp, u32, vp, vp, C.c_int, vp, vp, u64, u64, vp, vp, vp, C.c_int]
        L.or_splitmix64.restype = u64
        L.or_splitmix64.argtypes = [u64]
        L.or_go_sort.argtypes = [vp, vp, vp, C.c_int, C.c_int]
        L.or_choose.argtypes = [C.c_int, vp, vp, C.c_int, vp, C.c_int, C.c_int, vp, vp]
        L.or_rate.restype = i64
        L.or_rate.argtypes = [C.c_int, vp, C.c_int, C.c_double]
        _lib = L
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def go_sort(free, rl=None, rule=SORT_GO116):
    """sort.Sort on SortableGPUs; returns (sorted free, sorted rl, original indices)."""
    f = np.ascontiguousarray(free, dtype=np.int64).copy()
    r = np.zeros_like(f) if rl is None else np.ascontiguousarray(rl, dtype=np.int64).copy()
    idx = np.arange(len(f), dtype=np.int32)
    lib().or_go_sort(_p(f), _p(r), _p(idx), len(f), rule)
    return f, r, idx


def choose(policy, free, demand, rl=None, rule=SORT_GO116):
    """Rater.Choose on a bare GPU list -> (feasible, indexes, n_emitted)."""
    f = np.ascontiguousarray(free, dtype=np.int64)
    r = None if rl is None else np.ascontiguousarray(rl, dtype=np.int64)
    d = np.ascontiguousarray(demand, dtype=np.int32)
    out = np.full(max(len(d), 1), INT32_MIN, dtype=np.int32)
    ne = C.c_int32(0)
    ok = lib().or_choose(policy, _p(f), _p(r), len(f), _p(d), len(d), rule, _p(out), C.byref(ne))
    return bool(ok), out[: len(d)].tolist(), ne.value


def rate(policy, free, load_usage=0.0):
    f = np.ascontiguousarray(free, dtype=np.int64)
    return int(lib().or_rate(policy, _p(f), len(f), float(load_usage)))


def splitmix64(x):
    return int(lib().or_splitmix64(x & (2**64 - 1)))


class OracleCluster:
    """A cluster snapshot evaluated with the reference's algorithm on the CPU."""

    def __init__(self, n_nodes, sort_rule=SORT_GO116):
        self.n = int(n_nodes)
        self._h = lib().or_create(self.n, sort_rule)
        if not self._h:
            raise MemoryError("or_create failed")

    def close(self):
        if self._h:
            lib().or_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_buggy_restore(self, on):
        lib().or_set_buggy_restore(self._h, int(on))

    def node_upsert(self, idx, gpu_count):
        return lib().or_node_upsert(self._h, idx, gpu_count)

    def node_set(self, idx, free, rl=None):
        f = np.ascontiguousarray(free, dtype=np.int64)
        r = None if rl is None else np.ascontiguousarray(rl, dtype=np.int64)
        return lib().or_node_set(self._h, idx, len(f), _p(f), _p(r))

    def load_snapshot(self, gpu_count, free):
        """bulk: gpu_count [N] u8, free [N][8] int32 (columns >= gpu_count ignored)."""
        gc = np.ascontiguousarray(gpu_count, dtype=np.uint8)
        fr = np.ascontiguousarray(free, dtype=np.int32)
        if lib().or_snapshot_load(self._h, 0, len(gc), _p(gc), _p(fr)):
            raise ValueError("or_snapshot_load")

    def node_get(self, idx):
        g = C.c_int(0)
        f = np.zeros(MAX_GPUS, dtype=np.int64)
        r = np.zeros(MAX_GPUS, dtype=np.int64)
        rc = lib().or_node_get(self._h, idx, C.byref(g), _p(f), _p(r))
        if rc:
            raise IndexError(idx)
        return g.value, f[: g.value].copy(), r[: g.value].copy()

    def snapshot(self, width=8):
        assert width == 8
        free = np.zeros((self.n, 8), dtype=np.int32)
        rl = np.zeros((self.n, 8), dtype=np.int8)
        cnt = np.zeros(self.n, dtype=np.uint8)
        if lib().or_snapshot_read(self._h, 0, self.n, _p(cnt), _p(free), _p(rl)):
            raise ValueError("or_snapshot_read")
        return cnt, free, rl

    def node_apply(self, idx, gpu_idx, percent, sign):
        gi = np.ascontiguousarray(gpu_idx, dtype=np.int32)
        pc = np.ascontiguousarray(percent, dtype=np.int32)
        return lib().or_node_apply(self._h, idx, _p(gi), _p(pc), len(gi), sign)

    def load_add_node(self, idx, metric):
        return lib().or_load_add_node(self._h, idx, metric)

    def load_update(self, idx, metric, card, value, ts_sec, value_ok=True, ts_ok=True):
        return lib().or_load_update(self._h, idx, metric, card, float(value), int(value_ok), int(ts_sec), int(ts_ok))

    def load_update_bulk(self, metric, value, ts_sec, ok=None, first=0):
        """[n][8] cells; ok == 0 marks a cell whose strings did not parse"""
        v = np.ascontiguousarray(value, dtype=np.float64)
        t = np.ascontiguousarray(ts_sec, dtype=np.int64)
        k = None if ok is None else np.ascontiguousarray(ok, dtype=np.uint8)
        if lib().or_load_update_bulk(self._h, metric, first, v.shape[0], _p(v), _p(k), _p(t)):
            raise ValueError("or_load_update_bulk")

    def assume(self, percent, policy, idx=None, load=None, want_plan=False):
        d = np.ascontiguousarray(percent, dtype=np.int32)
        c = len(d)
        if idx is None:
            n, ip = self.n, None
        else:
            ia = np.ascontiguousarray(idx, dtype=np.uint32)
            n, ip = len(ia), _p(ia)
        feas = np.zeros(max(n, 1), dtype=np.uint8)
        reason = np.zeros(max(n, 1), dtype=np.int8)
        score = np.zeros(max(n, 1), dtype=np.int32)
        plan = np.full((max(n, 1), max(c, 1)), INT32_MIN, dtype=np.int32) if want_plan else None
        nem = np.zeros(max(n, 1), dtype=np.int32) if want_plan else None
        rc = lib().or_assume(self._h, ip, n, _p(d), c, policy, C.byref(load) if load is not None else None,
                             _p(feas), _p(reason), _p(score), _p(plan), _p(nem))
        if rc:
            raise ValueError("or_assume rc=%d" % rc)
        res = dict(feasible=feas[:n], reason=reason[:n], score=score[:n])
        if want_plan:
            res["plan"] = plan[:n, :c]
            res["n_emitted"] = nem[:n]
        return res

    def bind(self, idx, percent, policy, load=None):
        d = np.ascontiguousarray(percent, dtype=np.int32)
        out = np.full(max(len(d), 1), INT32_MIN, dtype=np.int32)
        sc = C.c_int32(0)
        rc = lib().or_bind(self._h, idx, _p(d), len(d), policy, C.byref(load) if load is not None else None,
                           _p(out), C.byref(sc))
        return rc, out[: len(d)].copy(), sc.value

    def schedule_batch(self, percent, n_containers, policy, load=None, seed=0, seq0=0, n_threads=1):
        pc = np.ascontiguousarray(percent, dtype=np.int32)
        P, max_c = pc.shape
        nc = np.ascontiguousarray(n_containers, dtype=np.int32)
        pol = np.ascontiguousarray(np.broadcast_to(np.asarray(policy, dtype=np.int32), (P,)))
        node = np.zeros(max(P, 1), dtype=np.int32)
        score = np.zeros(max(P, 1), dtype=np.int32)
        gidx = np.full((max(P, 1), max_c), INT32_MIN, dtype=np.int32)
        rc = lib().or_schedule_batch(self._h, P, _p(pc), _p(nc), max_c, _p(pol),
                                     C.byref(load) if load is not None else None, seed, seq0,
                                     _p(node), _p(score), _p(gidx), n_threads)
        if rc:
            raise RuntimeError("or_schedule_batch rc=%d" % rc)
        return node[:P], score[:P], gidx[:P]
