"""A Python restatement of Go 1.16 encoding/json's output rules — the checker the C++ encoder
(nano-gpu-scheduler_b200/host/extender_codec.cpp, bound by nano_gpu_scheduler_b200.routes) is compared with byte
for byte.  Go itself is not in this image, so this pins the codec against the published rules (encode.go:
escapeHTML on, \\u00xx for control bytes, \\ufffd for invalid UTF-8, U+2028/9 escaped, map keys sorted), not
against a Go run."""
from nano_gpu_scheduler_b200.routes import (HOST_LIB, Routes, _b, _take, decode_args, decode_binding, encode_binding_result,  # noqa: F401
                                            encode_filter, encode_filter_error, encode_priorities, json_string, lib, node_name,
                                            quantity)


# ---- Go 1.16 encoding/json, restated --------------------------------------------------------------
def go_json_string(s) -> bytes:
    """encode.go (*encodeState).string with escapeHTML = true, over the string's BYTES"""
    b = _b(s)
    out = bytearray(b'"')
    i = 0
    while i < len(b):
        c = b[i]
        if c < 0x80:
            if c == 0x22:
                out += b'\\"'
            elif c == 0x5C:
                out += b"\\\\"
            elif c == 0x0A:
                out += b"\\n"
            elif c == 0x0D:
                out += b"\\r"
            elif c == 0x09:
                out += b"\\t"
            elif c < 0x20 or c in (0x3C, 0x3E, 0x26):
                out += b"\\u00%02x" % c
            else:
                out.append(c)
            i += 1
            continue
        # utf8.DecodeRune: the longest valid sequence, else one byte -> U+FFFD
        for ln in (2, 3, 4):
            try:
                ch = b[i:i + ln].decode("utf-8")
            except UnicodeDecodeError:
                continue
            if len(ch) == 1:
                break
        else:
            out += b"\\ufffd"
            i += 1
            continue
        if ch in ("\u2028", "\u2029"):
            out += b"\\u%04x" % ord(ch)
        else:
            out += b[i:i + ln]
        i += ln
    out += b'"'
    return bytes(out)


def go_filter_result(can, failed: dict) -> bytes:
    """json.Marshal(&ExtenderFilterResult{NodeNames: &can, FailedNodes: failed, Error: ""}) (extender/v1 v0.18.0: no json tags)"""
    keys = sorted(failed, key=_b)
    return (b'{"Nodes":null,"NodeNames":[' + b",".join(go_json_string(x) for x in can) + b'],"FailedNodes":{'
            + b",".join(go_json_string(k) + b":" + go_json_string(failed[k]) for k in keys) + b'},"Error":""}')


def go_filter_error(msg) -> bytes:
    return b'{"Nodes":null,"NodeNames":null,"FailedNodes":null,"Error":' + go_json_string(msg) + b"}"


def go_priorities(hosts, scores) -> bytes:
    return b"[" + b",".join(b'{"Host":' + go_json_string(h) + b',"Score":%d}' % s for h, s in zip(hosts, scores)) + b"]"


def pod_json(percents, name="p", ns="default", uid="u1", quoted=True, extra=None):
    """the v1.Pod the kube-scheduler posts, reduced to what the dealer reads plus some fields it must skip"""
    cs = []
    for k, pc in enumerate(percents):
        c = {"name": "c%d" % k, "image": "busybox", "resources": {"requests": {"cpu": "100m"}}}
        if pc is not None:
            c["resources"]["limits"] = {"cpu": "1", "nano-gpu/gpu-percent": str(pc) if quoted else pc}
        cs.append(c)
    pod = {"kind": "Pod", "metadata": {"name": name, "namespace": ns, "uid": uid, "creationTimestamp": None, "labels": {"app": "x"}},
           "spec": {"containers": cs, "schedulerName": "default-scheduler", "tolerations": [{"key": "a", "effect": "NoSchedule"}]},
           "status": {"phase": "Pending"}}
    if extra:
        pod.update(extra)
    return pod
