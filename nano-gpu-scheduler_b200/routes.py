"""ctypes access to the C++ host layer above the C ABI (nano-gpu-scheduler_b200/host): the extender wire
codec and the filter / prioritize route bodies (pkg/routes/routes.go:40-123,
pkg/scheduler/predicate.go:21-38, priority.go:21-37 of the reference).  The flat C surface it binds is
host/extender_codec_c.cpp; nodes of a `Routes` cluster are named "node-%06d"."""
import ctypes as C
import json
import os
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
HOST_LIB = os.path.join(HERE, "libngs_dealer_host.so")

_lib = None


def lib():
    """the host library; built by nano_gpu_scheduler_b200.build.build_host()"""
    global _lib
    if _lib is None:
        L = C.CDLL(HOST_LIB)
        for name in ("ngsx_decode_args", "ngsx_decode_binding", "ngsx_json_string", "ngsx_encode_filter", "ngsx_encode_filter_error",
                     "ngsx_encode_binding_result", "ngsx_encode_priorities", "ngsx_routes_predicate", "ngsx_routes_prioritize",
                     "ngsx_routes_status", "ngsx_routes_bind_route", "ngsx_parse_policy"):
            getattr(L, name).restype = C.c_void_p
        L.ngsx_routes_new.restype = C.c_void_p
        L.ngsx_free.argtypes = [C.c_void_p]
        _lib = L
    return _lib

def _take(ptr, n):
    out = C.string_at(ptr, n.value)
    lib().ngsx_free(ptr)
    return out


def _b(s):
    return s if isinstance(s, bytes) else s.encode("utf-8", "surrogatepass")


def decode_args(body):
    body = _b(body)
    n = C.c_size_t()
    return json.loads(_take(lib().ngsx_decode_args(body, C.c_size_t(len(body)), C.byref(n)), n))


def decode_binding(body):
    body = _b(body)
    n = C.c_size_t()
    return json.loads(_take(lib().ngsx_decode_binding(body, C.c_size_t(len(body)), C.byref(n)), n))


def json_string(s):
    s = _b(s)
    n = C.c_size_t()
    return _take(lib().ngsx_json_string(s, C.c_size_t(len(s)), C.byref(n)), n)


def quantity(s):
    s = _b(s)
    v = C.c_int64()
    ok = lib().ngsx_quantity(s, C.c_size_t(len(s)), C.byref(v))
    return v.value if ok else None


def _blob(strings):
    bs = [_b(s) for s in strings]
    lens = (C.c_uint32 * max(len(bs), 1))(*[len(x) for x in bs])
    return b"".join(bs), lens


def encode_filter(can, failed):
    flat = list(can)
    for k, m in failed:
        flat += [k, m]
    blob, lens = _blob(flat)
    n = C.c_size_t()
    return _take(lib().ngsx_encode_filter(blob, lens, len(can), len(failed), C.byref(n)), n)


def encode_filter_error(msg):
    msg = _b(msg)
    n = C.c_size_t()
    return _take(lib().ngsx_encode_filter_error(msg, C.c_size_t(len(msg)), C.byref(n)), n)


def encode_binding_result(msg):
    msg = _b(msg)
    n = C.c_size_t()
    return _take(lib().ngsx_encode_binding_result(msg, C.c_size_t(len(msg)), C.byref(n)), n)


def encode_priorities(hosts, scores):
    blob, lens = _blob(hosts)
    sc = (C.c_int32 * max(len(scores), 1))(*scores)
    n = C.c_size_t()
    return _take(lib().ngsx_encode_priorities(blob, lens, sc, len(hosts), C.byref(n)), n)


def parse_policy(text):
    """spec.syncPeriod of a policy file -> [(name, period_ns)]; raises ValueError with the decoder's message"""
    b = _b(text)
    n = C.c_size_t()
    out = _take(lib().ngsx_parse_policy(b, C.c_size_t(len(b)), C.byref(n)), n).decode()
    if out.startswith("!"):
        raise ValueError(out[1:])
    return [(kv.rsplit("=", 1)[0], int(kv.rsplit("=", 1)[1])) for kv in out.split(";")] if out else []


def parse_duration(s):
    b = _b(s)
    v = C.c_int64()
    return v.value if lib().ngsx_parse_duration(b, C.c_size_t(len(b)), C.byref(v)) else None


def node_name(i):
    return "node-%06d" % i


class Routes:
    """The two extender routes over a device-resident dealer: JSON request body in, JSON answer out
    (ExtenderRoutes::Predicate / Prioritize of host/extender_codec.h).  gcount[i] = GPUs of node "node-%06d" % i;
    rater 0 binpack, 1 spread, 2 sample; assumed = [(node, uid, percent[], gpu_idx[])] pods that already run
    (NewDealer rebuilds their nodes, dealer.go:58-72).  Needs a CUDA device: there is no CPU fallback."""

    def __init__(self, gcount, rater, assumed=(), load=False, device=0):
        self.L = lib()
        g = np.ascontiguousarray(gcount, dtype=np.uint8)
        self.h = C.c_void_p(self.L.ngsx_routes_new(len(g), g.ctypes.data_as(C.c_void_p)))
        for node, uid, percent, gpu_idx in assumed:
            pc = (C.c_int32 * len(percent))(*percent)
            gi = (C.c_int32 * len(percent))(*gpu_idx)
            self.L.ngsx_routes_add_assumed(self.h, node, uid.encode(), pc, gi, len(percent))
        err = C.c_char_p()
        t0 = time.perf_counter()
        rc = self.L.ngsx_routes_start(self.h, rater, len(g) + 8, device, int(load), C.byref(err))
        self.start_ms = (time.perf_counter() - t0) * 1e3  # NewDealer: list the assumed pods, rebuild their nodes
        if rc != 0:
            msg = (err.value or b"").decode()
            self.close()
            raise RuntimeError("dealer start failed: " + msg)

    def close(self):
        if self.h:
            self.L.ngsx_routes_destroy(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def predicate(self, body):
        body = body if isinstance(body, bytes) else body.encode()
        n = C.c_size_t()
        return _take(self.L.ngsx_routes_predicate(self.h, body, C.c_size_t(len(body)), C.byref(n)), n)

    def prioritize(self, body):
        body = body if isinstance(body, bytes) else body.encode()
        n = C.c_size_t()
        p = C.c_int()
        out = _take(self.L.ngsx_routes_prioritize(self.h, body, C.c_size_t(len(body)), C.byref(n), C.byref(p)), n)
        return out, bool(p.value)

    def add_pod(self, ns, name, uid, percent, completed=False):
        """a pending pod the bind route will find (stands in for the apiserver's answer to Pods(ns).Get(name))"""
        pc = (C.c_int32 * max(len(percent), 1))(*percent)
        self.L.ngsx_routes_add_pod(self.h, ns.encode(), name.encode(), uid.encode(), pc, len(percent), int(completed))

    def bind_route(self, body):
        """BindRoute: ExtenderBindingArgs JSON in -> (ExtenderBindingResult JSON, HTTP status)"""
        body = body if isinstance(body, bytes) else body.encode()
        n = C.c_size_t()
        st = C.c_int()
        return _take(self.L.ngsx_routes_bind_route(self.h, body, C.c_size_t(len(body)), C.byref(n), C.byref(st)), n), st.value

    def status(self):
        """the /status body: every node the dealer has built, its GPUs' free percent and RemainLoad"""
        n = C.c_size_t()
        return _take(self.L.ngsx_routes_status(self.h, C.byref(n)), n)

    def counters(self):
        c = (C.c_uint64 * 7)()
        self.L.ngsx_routes_counters(self.h, c)
        return dict(requests=c[0], cache_hits=c[1], all_rows=c[2], ns_decode=c[3], ns_resolve=c[4], ns_dealer=c[5], ns_encode=c[6])

    def bind(self, node, uid, percent):
        pc = (C.c_int32 * len(percent))(*percent)
        out = (C.c_int32 * len(percent))()
        err = C.c_char_p()
        rc = self.L.ngsx_routes_bind(self.h, node.encode(), uid.encode(), pc, len(percent), out, C.byref(err))
        return rc, list(out), err.value
