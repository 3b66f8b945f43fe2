"""ctypes declarations of include/ngs_dealer.h — the same symbols the reference-side cgo stub
binds.  Loading fails loudly when the CUDA library has not been built: there
is no CPU fallback."""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("NGS_DEALER_LIB") or os.path.join(HERE, "libngs_dealer.so")  # override: profiling builds

MAX_GPUS = 8
MAX_GPUS_WIDE = 12
MAX_CONTAINERS = 12
MAX_PERIODS = 8
MAX_FREE = 65535
MAX_BATCH = 65536
INT32_MIN = -(2**31)

OK, ERR_INVALID, ERR_CUDA, ERR_UNKNOWN_NODE, ERR_INFEASIBLE, ERR_APPLY, ERR_RANGE, ERR_UNSUPPORTED, ERR_TIMEOUT = range(9)
POLICY_BINPACK, POLICY_SPREAD, POLICY_RANDOM = 0, 1, 2
REASON_OK, REASON_UNKNOWN_NODE, REASON_INFEASIBLE = 0, 1, 2
METRIC_CORE, METRIC_MEMORY = 0, 1

# every symbol include/ngs_dealer.h declares (tests check the library exports each one)
SYMBOLS = [
    "ngs_abi_version", "ngs_strerror", "ngs_last_error", "ngs_create", "ngs_destroy", "ngs_stream", "ngs_sync",
    "ngs_stats", "ngs_node_upsert", "ngs_snapshot_load", "ngs_snapshot_read", "ngs_node_apply", "ngs_release_batch", "ngs_load_add_node",
    "ngs_load_update", "ngs_load_update_bulk", "ngs_parse_usage", "ngs_parse_update_time", "ngs_assume", "ngs_assume_dev", "ngs_score",
    "ngs_bind", "ngs_schedule_batch", "ngs_schedule_batch_dev", "ngs_xchg_export", "ngs_xchg_attach",
    "ngs_schedule_batch_noalloc", "ngs_node_set_wide", "ngs_node_get_wide", "ngs_score_view", "ngs_serve_pod", "ngs_serve_stop", "ngs_serve_idle_ms", "ngs_serve_verbs", "ngs_assume_failed_plans",
]
NOALLOC_MAX_PODS = 256
NOALLOC_MAX_CTAS = 160
ERR_NOALLOC_LIMIT = 9
NODE_NONE, NODE_INVALID_POD = -1, -2
MAX_RANKS = 8
IPC_HANDLE_BYTES = 64


class LoadCtx(C.Structure):
    _fields_ = [
        ("n_periods", C.c_int32),
        ("metric", C.c_int32 * MAX_PERIODS),
        ("window_ns", C.c_int64 * MAX_PERIODS),
        ("now_ns", C.c_int64),
    ]


class SnapshotView(C.Structure):
    """ngs_snapshot_view: caller-owned snapshot tables for ngs_schedule_batch_noalloc"""
    _fields_ = [
        ("free", C.c_void_p), ("gpu_count", C.c_void_p), ("remain_load", C.c_void_p), ("tenths", C.c_void_p),
        ("ts", C.c_void_p), ("plan_tag", C.c_void_p), ("plan_score", C.c_void_p),
        ("n_nodes", C.c_uint32), ("first_node", C.c_uint32),
    ]


class Stats(C.Structure):
    _fields_ = [
        ("node_evals", C.c_uint64),
        ("kernel_launches", C.c_uint64),
        ("pods_scheduled", C.c_uint64),
        ("h2d_bytes", C.c_uint64),
        ("d2h_bytes", C.c_uint64),
    ]


_lib = None


def load():
    """dlopen the in-tree CUDA library and declare the prototypes."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). The dealer has no CPU fallback."
        )
    L = C.CDLL(LIB_PATH)
    vp, u32, i32, i64, u64, ci, dbl = C.c_void_p, C.c_uint32, C.c_int32, C.c_int64, C.c_uint64, C.c_int, C.c_double
    L.ngs_abi_version.restype = u32
    L.ngs_abi_version.argtypes = []
    L.ngs_strerror.restype = C.c_char_p
    L.ngs_strerror.argtypes = [ci]
    L.ngs_last_error.restype = C.c_char_p
    L.ngs_last_error.argtypes = [vp]
    L.ngs_create.argtypes = [C.POINTER(vp), ci, u32, u32, vp]
    L.ngs_destroy.argtypes = [vp]
    L.ngs_stream.restype = vp
    L.ngs_stream.argtypes = [vp]
    L.ngs_sync.argtypes = [vp]
    L.ngs_stats.argtypes = [vp, C.POINTER(Stats)]
    L.ngs_node_upsert.argtypes = [vp, u32, ci]
    L.ngs_node_set_wide.argtypes = [vp, u32, ci, vp, vp]
    L.ngs_node_get_wide.argtypes = [vp, u32, C.POINTER(ci), vp, vp]
    L.ngs_snapshot_load.argtypes = [vp, u32, u32, vp, vp, vp]
    L.ngs_snapshot_read.argtypes = [vp, u32, u32, vp, vp, vp]
    L.ngs_node_apply.argtypes = [vp, u32, vp, vp, ci, ci]
    L.ngs_release_batch.argtypes = [vp, u32, vp, vp, vp, vp, ci, vp]
    L.ngs_load_add_node.argtypes = [vp, u32, ci]
    L.ngs_load_update.argtypes = [vp, u32, ci, ci, dbl, ci, i64, ci]
    L.ngs_load_update_bulk.argtypes = [vp, ci, u32, u32, vp, vp, vp, vp]
    L.ngs_parse_usage.argtypes = [C.c_char_p, C.POINTER(dbl)]
    L.ngs_parse_update_time.argtypes = [C.c_char_p, C.POINTER(i64)]
    L.ngs_assume.argtypes = [vp, vp, u32, vp, ci, ci, vp, vp, vp, vp, vp, vp]
    L.ngs_assume_failed_plans.argtypes = [vp, vp, u32, vp, ci, ci, vp, vp, vp, vp, vp, vp]
    L.ngs_assume_dev.argtypes = [vp, u32, vp, ci, ci, vp, vp, vp]
    L.ngs_score.argtypes = [vp, vp, u32, vp, ci, ci, vp, vp]
    L.ngs_score_view.argtypes = [vp, vp, u32, vp, ci, ci, vp, C.POINTER(vp)]
    L.ngs_bind.argtypes = [vp, u32, vp, ci, ci, vp, vp, vp]
    L.ngs_serve_pod.argtypes = [vp, vp, ci, ci, ci, vp, u64, u64, vp, vp, vp]
    L.ngs_serve_stop.argtypes = [vp]
    L.ngs_serve_idle_ms.argtypes = [vp, dbl]
    L.ngs_serve_verbs.argtypes = [vp, ci]
    L.ngs_schedule_batch.argtypes = [vp, u32, vp, vp, ci, vp, vp, u64, u64, vp, vp, vp]
    L.ngs_schedule_batch_dev.argtypes = [vp, u32, vp, vp, ci, vp, vp, u64, u64, vp, vp, vp]
    L.ngs_xchg_export.argtypes = [vp, vp]
    L.ngs_xchg_attach.argtypes = [vp, ci, ci, vp]
    L.ngs_schedule_batch_noalloc.argtypes = [ci, vp, C.POINTER(SnapshotView), u32, vp, vp, ci, vp, vp, u64, u64, vp, vp, vp]
    for name in SYMBOLS:
        fn = getattr(L, name)
        if fn.restype is C.c_int and name not in ("ngs_parse_usage", "ngs_parse_update_time"):
            fn.restype = ci
    _lib = L
    return L
