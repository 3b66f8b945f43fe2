"""numpy-facing wrapper of the C ABI (include/ngs_dealer.h).  Thin by design: every method is one
C call; no evaluation logic lives in Python and nothing here falls back to the CPU."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _abi
from ._abi import (INT32_MIN, MAX_CONTAINERS, MAX_GPUS, MAX_GPUS_WIDE, METRIC_CORE, METRIC_MEMORY, POLICY_BINPACK,  # noqa: F401
                   POLICY_RANDOM, POLICY_SPREAD, LoadCtx, Stats)


class NgsError(RuntimeError):
    def __init__(self, status, message):
        super().__init__(f"[{status}] {message}")
        self.status = status


def make_load_ctx(periods, now_ns):
    """periods: [(metric, window_ns)] — PolicySpec.SyncPeriod after getActiveDuration
    (/root/reference/pkg/dealer/stats.go:57-66); now_ns: the evaluation's time.Now()."""
    L = LoadCtx()
    L.n_periods = len(periods)
    for i, (m, w) in enumerate(periods[: _abi.MAX_PERIODS]):
        L.metric[i] = m
        L.window_ns[i] = w
    L.now_ns = now_ns
    return L


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _ref(load):
    return None if load is None else C.c_void_p(C.addressof(load))


class GpuDealer:
    """One snapshot shard resident on one GPU."""

    def __init__(self, max_nodes, device=0, first_node=0, stream=None):
        self._lib = _abi.load()
        self._h = C.c_void_p()
        self.n = int(max_nodes)
        self.first = int(first_node)
        rc = self._lib.ngs_create(C.byref(self._h), device, self.n, self.first, stream)
        if rc != _abi.OK:
            msg = self._lib.ngs_last_error(self._h).decode() if self._h else ""
            if self._h:
                self._lib.ngs_destroy(self._h)
                self._h = C.c_void_p()
            raise NgsError(rc, msg or self._lib.ngs_strerror(rc).decode())

    # -- lifetime ---------------------------------------------------------------------------
    def close(self):
        if getattr(self, "_h", None):
            self._lib.ngs_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def _check(self, rc):
        if rc != _abi.OK:
            raise NgsError(rc, self._lib.ngs_last_error(self._h).decode() or self._lib.ngs_strerror(rc).decode())

    @property
    def stream(self):
        return self._lib.ngs_stream(self._h)

    def sync(self):
        self._check(self._lib.ngs_sync(self._h))

    def stats(self):
        s = Stats()
        self._check(self._lib.ngs_stats(self._h, C.byref(s)))
        return {k: int(getattr(s, k)) for k, _ in Stats._fields_}

    # -- snapshot ---------------------------------------------------------------------------
    def node_upsert(self, idx, gpu_count):
        self._check(self._lib.ngs_node_upsert(self._h, idx, gpu_count))

    def node_set_wide(self, idx, gpu_count, free=None, remain_load=None):
        """one node of any width (up to MAX_GPUS_WIDE GPUs) with all its cells"""
        fr = None if free is None else np.ascontiguousarray(free, dtype=np.int32)
        rl = None if remain_load is None else np.ascontiguousarray(remain_load, dtype=np.int8)
        for a in (fr, rl):
            if a is not None and len(a) < gpu_count:
                raise ValueError("node_set_wide: fewer cells than gpu_count")
        self._check(self._lib.ngs_node_set_wide(self._h, idx, gpu_count, _p(fr), _p(rl)))

    def node_get_wide(self, idx):
        g = C.c_int(0)
        fr = np.zeros(MAX_GPUS_WIDE, dtype=np.int32)
        rl = np.zeros(MAX_GPUS_WIDE, dtype=np.int8)
        self._check(self._lib.ngs_node_get_wide(self._h, idx, C.byref(g), _p(fr), _p(rl)))
        return g.value, fr, rl

    def snapshot_load(self, gpu_count, free=None, remain_load=None, first=None):
        gc = np.ascontiguousarray(gpu_count, dtype=np.uint8)
        fr = None if free is None else np.ascontiguousarray(free, dtype=np.int32)
        rl = None if remain_load is None else np.ascontiguousarray(remain_load, dtype=np.int8)
        first = self.first if first is None else first
        self._check(self._lib.ngs_snapshot_load(self._h, first, len(gc), _p(gc), _p(fr), _p(rl)))

    def snapshot_read(self, first=None, n=None):
        first = self.first if first is None else first
        n = self.n if n is None else n
        gc = np.zeros(n, dtype=np.uint8)
        fr = np.zeros((n, MAX_GPUS), dtype=np.int32)
        rl = np.zeros((n, MAX_GPUS), dtype=np.int8)
        self._check(self._lib.ngs_snapshot_read(self._h, first, n, _p(gc), _p(fr), _p(rl)))
        return gc, fr, rl

    def node_apply(self, idx, gpu_idx, percent, sign):
        gi = np.ascontiguousarray(gpu_idx, dtype=np.int32)
        pc = np.ascontiguousarray(percent, dtype=np.int32)
        return self._lib.ngs_node_apply(self._h, idx, _p(gi), _p(pc), len(pc), sign)

    def release_batch(self, node, gpu_idx, percent, n_containers):
        """GPUs.Release for many pods at once -> (status of the call, per-pod status)"""
        nd = np.ascontiguousarray(node, dtype=np.uint32)
        gi = np.ascontiguousarray(gpu_idx, dtype=np.int32)
        pc = np.ascontiguousarray(percent, dtype=np.int32)
        nc = np.ascontiguousarray(n_containers, dtype=np.int32)
        n = len(nd)
        max_c = gi.shape[1] if gi.ndim == 2 else 1
        st = np.zeros(max(n, 1), dtype=np.int32)
        rc = self._lib.ngs_release_batch(self._h, n, _p(nd), _p(gi), _p(pc), _p(nc), max_c, _p(st))
        return rc, st[:n]

    # -- load tables ------------------------------------------------------------------------
    def load_add_node(self, idx, metric):
        self._check(self._lib.ngs_load_add_node(self._h, idx, metric))

    def load_update(self, idx, metric, card, value, ts_sec, value_ok=True, ts_ok=True):
        self._check(self._lib.ngs_load_update(self._h, idx, metric, card, float(value), int(value_ok), int(ts_sec), int(ts_ok)))

    def load_update_bulk(self, metric, value, ts_sec, value_ok=None, ts_ok=None, first=None):
        v = np.ascontiguousarray(value, dtype=np.float64)
        t = np.ascontiguousarray(ts_sec, dtype=np.int64)
        vo = None if value_ok is None else np.ascontiguousarray(value_ok, dtype=np.uint8)
        to = None if ts_ok is None else np.ascontiguousarray(ts_ok, dtype=np.uint8)
        first = self.first if first is None else first
        self._check(self._lib.ngs_load_update_bulk(self._h, metric, first, v.shape[0], _p(v), _p(vo), _p(t), _p(to)))

    # -- verbs ------------------------------------------------------------------------------
    def assume(self, percent, policy, idx=None, load=None, want_plan=False, n=None, failed_plans_only=False):
        """failed_plans_only (with want_plan): plans / emitted counts of the infeasible candidates only (ngs_assume_failed_plans)"""
        d = np.ascontiguousarray(percent, dtype=np.int32)
        c = len(d)
        if idx is None:
            n, ia = (self.n if n is None else n), None
        else:
            ia = np.ascontiguousarray(idx, dtype=np.uint32)
            n = len(ia)
        feas = np.zeros(max(n, 1), dtype=np.uint8)
        reason = np.zeros(max(n, 1), dtype=np.int8)
        score = np.zeros(max(n, 1), dtype=np.int32)
        plan = np.full((max(n, 1), max(c, 1)), INT32_MIN, dtype=np.int32) if want_plan else None
        nem = np.zeros(max(n, 1), dtype=np.int32) if want_plan else None
        fn = self._lib.ngs_assume_failed_plans if (want_plan and failed_plans_only) else self._lib.ngs_assume
        self._check(fn(self._h, _p(ia), n, _p(d), c, policy, _ref(load), _p(feas), _p(reason), _p(score), _p(plan), _p(nem)))
        res = dict(feasible=feas[:n], reason=reason[:n], score=score[:n])
        if want_plan:
            res["plan"] = plan[:n, :c]
            res["n_emitted"] = nem[:n]
        return res

    def assume_dev(self, percent, policy, d_feasible, d_score, load=None, n=None):
        """filter + score rows 0..n-1, outputs to raw device addresses, no synchronisation"""
        d = np.ascontiguousarray(percent, dtype=np.int32)
        self._check(self._lib.ngs_assume_dev(self._h, self.n if n is None else n, _p(d), len(d), policy, _ref(load),
                                             d_feasible, d_score))

    def score(self, percent, policy, idx=None, load=None, n=None):
        d = np.ascontiguousarray(percent, dtype=np.int32)
        if idx is None:
            n, ia = (self.n if n is None else n), None
        else:
            ia = np.ascontiguousarray(idx, dtype=np.uint32)
            n = len(ia)
        score = np.zeros(max(n, 1), dtype=np.int32)
        self._check(self._lib.ngs_score(self._h, _p(ia), n, _p(d), len(d), policy, _ref(load), _p(score)))
        return score[:n]

    def score_view(self, percent, policy, idx=None, load=None, n=None):
        """score() without the copy: a read-only array over the handle's own pinned memory, valid until the next call"""
        d = np.ascontiguousarray(percent, dtype=np.int32)
        if idx is None:
            n, ia = (self.n if n is None else n), None
        else:
            ia = np.ascontiguousarray(idx, dtype=np.uint32)
            n = len(ia)
        ptr = C.c_void_p(0)
        self._check(self._lib.ngs_score_view(self._h, _p(ia), n, _p(d), len(d), policy, _ref(load), C.byref(ptr)))
        if n == 0:
            return np.zeros(0, dtype=np.int32)
        a = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_int32)), shape=(n,))
        a.flags.writeable = False
        return a

    def bind(self, idx, percent, policy, load=None):
        """-> (status, gpu_idx[C], score); status != 0 leaves the snapshot untouched."""
        d = np.ascontiguousarray(percent, dtype=np.int32)
        out = np.full(max(len(d), 1), INT32_MIN, dtype=np.int32)
        sc = C.c_int32(0)
        rc = self._lib.ngs_bind(self._h, idx, _p(d), len(d), policy, _ref(load), _p(out), C.byref(sc))
        return rc, out[: len(d)].copy(), sc.value

    def schedule_batch(self, percent, n_containers, policy, load=None, seed=0, seq0=0):
        pc = np.ascontiguousarray(percent, dtype=np.int32)
        P, max_c = pc.shape
        nc = np.ascontiguousarray(n_containers, dtype=np.int32)
        pol = np.ascontiguousarray(np.broadcast_to(np.asarray(policy, dtype=np.int32), (P,)))
        node = np.zeros(max(P, 1), dtype=np.int32)
        score = np.zeros(max(P, 1), dtype=np.int32)
        gidx = np.full((max(P, 1), max_c), INT32_MIN, dtype=np.int32)
        self._check(self._lib.ngs_schedule_batch(self._h, P, _p(pc), _p(nc), max_c, _p(pol), _ref(load), seed, seq0,
                                                 _p(node), _p(score), _p(gidx)))
        return node[:P], score[:P], gidx[:P]

    def serve_pod(self, percent, policy, load=None, seed=0, seq0=0):
        """one pod through the resident kernel -> (node, score, gpu_idx[C]); same decision as schedule_batch with one pod"""
        d = np.ascontiguousarray(percent, dtype=np.int32)
        c = max(len(d), 1)
        node, score = C.c_int32(0), C.c_int32(0)
        gidx = np.full(c, INT32_MIN, dtype=np.int32)
        self._check(self._lib.ngs_serve_pod(self._h, _p(d), len(d), c, policy, _ref(load), seed, seq0, C.byref(node), C.byref(score), _p(gidx)))
        return node.value, score.value, gidx[: len(d)]

    def serve_stop(self):
        self._check(self._lib.ngs_serve_stop(self._h))

    def serve_verbs(self, on=True):
        """assume / score / bind through the resident kernel as well"""
        self._check(self._lib.ngs_serve_verbs(self._h, int(bool(on))))

    def serve_idle_ms(self, ms):
        self._check(self._lib.ngs_serve_idle_ms(self._h, float(ms)))

    def schedule_batch_dev(self, n_pods, d_percent, d_ncont, max_c, d_policy, d_node, d_score, d_gidx, load=None, seed=0,
                           seq0=0):
        """device-pointer form: arguments are raw device addresses (e.g. torch.Tensor.data_ptr())."""
        self._check(self._lib.ngs_schedule_batch_dev(self._h, n_pods, d_percent, d_ncont, max_c, d_policy, _ref(load), seed,
                                                     seq0, d_node, d_score, d_gidx))


    # -- node-range sharding over the GPUs of one box ------------------------------------------
    def xchg_export(self) -> bytes:
        buf = C.create_string_buffer(_abi.IPC_HANDLE_BYTES)
        self._check(self._lib.ngs_xchg_export(self._h, buf))
        return buf.raw

    def xchg_attach(self, rank, world, handles: bytes):
        assert len(handles) == world * _abi.IPC_HANDLE_BYTES
        self._check(self._lib.ngs_xchg_attach(self._h, rank, world, handles))


def schedule_batch_noalloc(view, n_pods, d_percent, d_ncont, max_c, d_policy, d_node, d_score, d_gidx, load=None, seed=0,
                           seq0=0, device=0, stream=None):
    """ngs_schedule_batch_noalloc: schedule a batch on caller-owned tables without any allocation by
    the library (hosts whose device memory belongs to another allocator, or that must stay clear of
    cudaMalloc's device-wide synchronisation on the scheduling path).  `view` is an
    _abi.SnapshotView of device pointers, the other pointer arguments are raw device addresses.
    At most _abi.NOALLOC_MAX_PODS pods; waits for the batch and returns its status code.
    Calls on one device must not overlap: they share one fixed scratch array."""
    return _abi.load().ngs_schedule_batch_noalloc(device, stream, C.byref(view), n_pods, d_percent, d_ncont, max_c, d_policy,
                                                  _ref(load), seed, seq0, d_node, d_score, d_gidx)


def parse_usage(s: str):
    v = C.c_double(0.0)
    ok = _abi.load().ngs_parse_usage(s.encode(), C.byref(v))
    return bool(ok), v.value


def parse_update_time(s: str):
    t = C.c_int64(0)
    ok = _abi.load().ngs_parse_update_time(s.encode(), C.byref(t))
    return bool(ok), t.value
