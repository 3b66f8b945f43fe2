"""The extender wire codec on the CPU: decode of ExtenderArgs / ExtenderBindingArgs, the
Quantity grammar, and byte-exact encoding of ExtenderFilterResult / HostPriorityList against a Python
restatement of Go 1.16 encoding/json (tests/extender_util.py).  Reference call sites:
pkg/routes/routes.go:47-53,75-87,113-119,139-150; pkg/scheduler/predicate.go:21-38; priority.go:30-37."""
import ctypes as C
import json
import random

import pytest

from tests import extender_util as x


@pytest.fixture(scope="module", autouse=True)
def _built(built):
    import nano_gpu_scheduler_b200.build as b

    b.build_host()


def test_c_surface_exports():
    L = C.CDLL(x.HOST_LIB)
    for name in ("ngsx_decode_args", "ngsx_decode_binding", "ngsx_json_string", "ngsx_quantity", "ngsx_encode_filter",
                 "ngsx_encode_filter_error", "ngsx_encode_binding_result", "ngsx_encode_priorities", "ngsx_node_index_check",
                 "ngsx_routes_new", "ngsx_routes_add_assumed", "ngsx_routes_start", "ngsx_routes_predicate", "ngsx_routes_prioritize",
                 "ngsx_routes_counters", "ngsx_routes_bind", "ngsx_routes_destroy", "ngsx_free", "ngsx_routes_status", "ngsx_routes_add_pod",
                 "ngsx_routes_bind_route", "ngsx_parse_policy", "ngsx_parse_duration"):
        assert hasattr(L, name), name


STRINGS = [
    "", "node-1", "a\"b", "back\\slash", "tab\tnl\ncr\r", "\x00\x01\x1f", "\x08\x0c", "<html>&amp;</html>", "\x7f",
    "ünïcödé", "日本語ノード", "emoji 😀", "line\u2028sep\u2029", b"bad \xff\xfe utf8", b"trunc \xe2\x82", b"overlong \xc0\xaf",
    b"surrogate \xed\xa0\x80", b"\xf4\x90\x80\x80 beyond", "can't allocate   [(20) (40)] on (10)(50): indexes=[1]",
    'node "node-000009" not found',
]


def test_string_encoding_matches_go_rules():
    for s in STRINGS:
        got = x.json_string(s)
        assert got == x.go_json_string(s), (s, got)
        if isinstance(s, str):
            assert json.loads(got) == s  # and it is JSON that reads back
    rng = random.Random(7)
    for _ in range(400):
        raw = bytes(rng.choice([rng.randrange(256), rng.randrange(0x20, 0x7F)]) for _ in range(rng.randrange(0, 40)))
        assert x.json_string(raw) == x.go_json_string(raw), raw


def test_filter_result_bytes():
    can = ["node-b", "node-a", "node-a"]  # order of arrival, duplicates kept (predicate.go:29)
    failed = {"z<1>": "can't allocate   [(20)] on (10): indexes=[]", "a&b": 'nano gpu scheduler get node failed: node "a&b" not found', "m": "x"}
    got = x.encode_filter(can, sorted(failed.items()))
    assert got == x.go_filter_result(can, failed)
    back = json.loads(got)
    assert back == {"Nodes": None, "NodeNames": can, "FailedNodes": failed, "Error": ""}
    # nothing fits / everything fits
    assert x.encode_filter([], []) == b'{"Nodes":null,"NodeNames":[],"FailedNodes":{},"Error":""}'
    assert x.encode_filter_error("nano-gpu-scheduler extender must be configured with nodeCacheCapable=true") == x.go_filter_error(
        "nano-gpu-scheduler extender must be configured with nodeCacheCapable=true")
    assert x.encode_filter_error('invalid character \'x\' looking for beginning of value') == \
        b'{"Nodes":null,"NodeNames":null,"FailedNodes":null,"Error":"invalid character \'x\' looking for beginning of value"}'


def test_priority_list_and_binding_result_bytes():
    hosts = ["n1", "n<2>", "日本"]
    scores = [0, 92, -8]
    got = x.encode_priorities(hosts, scores)
    assert got == x.go_priorities(hosts, scores)
    assert json.loads(got) == [{"Host": h, "Score": s} for h, s in zip(hosts, scores)]
    assert x.encode_priorities([], []) == b"[]"
    assert x.encode_binding_result("") == b'{"Error":""}'
    assert x.encode_binding_result('pods "p" not found') == b'{"Error":"pods \\"p\\" not found"}'


def test_decode_extender_args():
    names = ["node-%06d" % i for i in range(2000)]
    body = json.dumps({"Pod": x.pod_json([20, None, 0, 100]), "Nodes": None, "NodeNames": names})
    d = x.decode_args(body)
    assert d["err"] == "" and d["has_pod"] and d["has_node_names"] and not d["has_nodes"]
    assert d["node_names"] == names
    assert d["pod"]["name"] == "p" and d["pod"]["namespace"] == "default" and d["pod"]["uid"] == "u1"
    assert d["pod"]["labels"] == {"app": "x"} and d["pod"]["annotations"] == {}
    assert d["pod"]["containers"] == [
        {"name": "c0", "has": True, "percent": 20}, {"name": "c1", "has": False, "percent": 0},
        {"name": "c2", "has": True, "percent": 0}, {"name": "c3", "has": True, "percent": 100}]
    # indent, unquoted quantities, keys in another case, unknown members, escapes in names and annotations
    pod = x.pod_json([35], quoted=False)
    pod["metadata"]["annotations"] = {"nano-gpu/assume": "true", "nano-gpu/container-c0": "3", "note": "a\"b\\cé😀"}
    pod["spec"]["nodeName"] = "node-7"
    body = json.dumps({"unknown": [1, {"a": [None, True, 1.5e3]}], "pod": pod, "NODENAMES": ["a\tb", "plain", "ü", "x/y"]}, indent=2)
    d = x.decode_args(body)
    assert d["err"] == "" and d["node_names"] == ["a\tb", "plain", "ü", "x/y"]
    assert d["pod"]["containers"] == [{"name": "c0", "has": True, "percent": 35}]
    assert d["pod"]["annotations"]["note"] == "a\"b\\cé😀" and d["pod"]["nodeName"] == "node-7"
    # raw UTF-8 in the body (ensure_ascii off) reads the same
    d2 = x.decode_args(json.dumps({"Pod": pod, "NodeNames": ["ü", "日本"]}, ensure_ascii=False))
    assert d2["node_names"] == ["ü", "日本"] and d2["pod"]["annotations"]["note"] == "a\"b\\cé😀"
    # nodeCacheCapable=false shape: no NodeNames -> nil (routes.go:62 answers with an error)
    d = x.decode_args(json.dumps({"Pod": x.pod_json([10]), "Nodes": {"items": [{"metadata": {"name": "n"}}]}}))
    assert d["err"] == "" and not d["has_node_names"] and d["has_nodes"]
    assert not x.decode_args('{"Pod": null, "NodeNames": null}')["has_node_names"]
    d = x.decode_args('{"NodeNames": []}')
    assert d["has_node_names"] and d["node_names"] == [] and not d["has_pod"]
    # later duplicate wins, like decoding twice into the same struct
    assert x.decode_args('{"NodeNames": ["a"], "NodeNames": ["b", "c"]}')["node_names"] == ["b", "c"]
    # lone surrogate escapes and invalid UTF-8 become U+FFFD (decode.go unquote)
    assert x.decode_args(b'{"NodeNames": ["\\ud800x", "\\udc00", "a\xffb"]}')["node_names"] == ["�x", "�", "a�b"]


GO_ERRORS = [
    (b"", "EOF"),
    (b"   \n", "EOF"),
    (b"{", "unexpected EOF"),
    (b'{"NodeNames": ["a"', "unexpected EOF"),
    (b'{"NodeNames": ["a', "unexpected EOF"),
    (b"x", "invalid character 'x' looking for beginning of value"),
    (b'{"a":\x00}', "invalid character '\\x00' looking for beginning of value"),
    (b'{"a" 1}', "invalid character '1' after object key"),
    (b'{"a":1 "b":2}', "invalid character '\"' after object key:value pair"),
    (b'{"NodeNames":["a" "b"]}', "invalid character '\"' after array element"),
    (b'{"a":tru}', "invalid character '}' in literal true (expecting 'e')"),
    (b'{a:1}', "invalid character 'a' looking for beginning of object key string"),
    (b'{"NodeNames":["\\q"]}', "invalid character 'q' in string escape code"),
    (b'{"NodeNames":["\\u12G4"]}', "invalid character 'G' in \\u hexadecimal character escape"),
    (b'{"NodeNames":["a\nb"]}', "invalid character '\\n' in string literal"),
    (b'{"a":-}', "invalid character '}' in numeric literal"),
    (b'{"a":1.}', "invalid character '}' after decimal point in numeric literal"),
    (b'{"a":1e}', "invalid character '}' in exponent of numeric literal"),
    (b'{"NodeNames": [1]}', "json: cannot unmarshal number into Go struct field ExtenderArgs.NodeNames of type string"),
    (b'{"NodeNames": "x"}', "json: cannot unmarshal string into Go struct field ExtenderArgs.NodeNames of type []string"),
    (b'{"NodeNames": ["a", {"b": 1}, "c"]}', "json: cannot unmarshal object into Go struct field ExtenderArgs.NodeNames of type string"),
    (b"[1,2]", "json: cannot unmarshal array into Go value of type v1.ExtenderArgs"),
]


def test_decode_errors_use_go_texts():
    for body, want in GO_ERRORS:
        assert x.decode_args(body)["err"] == want, body
    # a type mismatch does not stop the decode: the rest of the document is still read (decode.go saveError)
    d = x.decode_args(b'{"NodeNames": ["a", 5, "c"], "Pod": {"metadata": {"name": "late"}}}')
    assert d["err"].startswith("json: cannot unmarshal number") and d["node_names"] == ["a", "", "c"] and d["pod"]["name"] == "late"
    # trailing bytes after the first value are the next document, not an error (json.Decoder)
    assert x.decode_args(b'{"NodeNames": ["a"]} trailing')["err"] == ""
    # a bad quantity fails the decode with apimachinery's message
    d = x.decode_args(json.dumps({"Pod": x.pod_json(["lots"]), "NodeNames": []}))
    assert d["err"].startswith("quantities must match the regular expression")


def test_quantity_values():
    table = {"20": 20, "100": 100, "0": 0, " 30 ": 30, "+5": 5, "0.5": 1, "1.0": 1, "2.000": 2, "500m": 1, "1500m": 2, "2000m": 2,
             "1k": 1000, "1M": 10**6, "1Ki": 1024, "3Mi": 3 * 2**20, "1e2": 100, "1E2": 100, "2.5e1": 25, "25e-1": 3, "1n": 1, "0m": 0,
             "12345678901234567890": 2**63 - 1, "9E": 9 * 10**18, "10E": 2**63 - 1}
    for s, want in table.items():
        assert x.quantity(s) == want, s
    for bad in ("", "abc", "1x", "1.2.3", "e5", "1e", "1Kii", "--1", "1 0"):
        assert x.quantity(bad) is None, bad


def test_decode_binding_args():
    d = x.decode_binding(json.dumps({"PodName": "p", "PodNamespace": "ns", "PodUID": "abc-123", "Node": "node-000004"}))
    assert d == {"err": "", "PodName": "p", "PodNamespace": "ns", "PodUID": "abc-123", "Node": "node-000004"}
    d = x.decode_binding(b'{"podname": "q", "extra": {}, "Node": null}')
    assert d["PodName"] == "q" and d["Node"] == "" and d["err"] == ""
    assert x.decode_binding(b'{"Node": 4}')["err"] == "json: cannot unmarshal number into Go struct field ExtenderBindingArgs.Node of type string"


def test_node_index():
    assert x.lib().ngsx_node_index_check(200000) == 0


def test_routes_fail_loudly_without_a_gpu():
    import numpy as np
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    L = x.lib()
    g = np.full(4, 8, dtype=np.uint8)
    h = C.c_void_p(L.ngsx_routes_new(4, g.ctypes.data_as(C.c_void_p)))
    err = C.c_char_p()
    assert L.ngsx_routes_start(h, 0, 16, 0, 0, C.byref(err)) == 1 and b"CUDA" in err.value  # no CPU fallback behind the routes
    L.ngsx_routes_destroy(h)


def test_decode_fuzz_against_python_json():
    """random well-formed bodies (any unicode, either escaping style, any indentation, extra members) read the same
    as Python's json reads them; random byte damage to such bodies never crashes the scanner"""
    from hypothesis import given, settings
    from hypothesis import strategies as st

    names = st.lists(st.text(max_size=24), max_size=40)
    junk = st.recursive(st.none() | st.booleans() | st.integers(-10**6, 10**6) | st.floats(allow_nan=False, allow_infinity=False) | st.text(max_size=8),
                        lambda ch: st.lists(ch, max_size=4) | st.dictionaries(st.text(max_size=6), ch, max_size=4), max_leaves=12)

    @settings(max_examples=150, deadline=None)
    @given(names, st.lists(st.one_of(st.none(), st.integers(0, 400)), max_size=6), st.dictionaries(st.text(max_size=10), st.text(max_size=10), max_size=5),
           junk, st.booleans(), st.sampled_from([None, 0, 3]))
    def run(node_names, percents, annotations, extra, ascii_only, indent):
        pod = x.pod_json(percents, name="pé", uid="uid-日")
        pod["metadata"]["annotations"] = annotations
        body = json.dumps({"zz": extra, "Pod": pod, "NodeNames": node_names, "aa": [extra]}, ensure_ascii=ascii_only, indent=indent)
        d = x.decode_args(body.encode("utf-8", "surrogatepass"))
        want_names = [s.encode("utf-8", "surrogatepass").decode("utf-8", "replace") if not ascii_only else s for s in node_names]
        if ascii_only:  # lone surrogates arrive as \\udXXX escapes: Go makes them U+FFFD
            want_names = ["".join("�" if 0xD800 <= ord(c) < 0xE000 else c for c in s) for s in node_names]
        assert d["err"] == ""
        if not ascii_only and any(0xD800 <= ord(c) < 0xE000 for s in node_names for c in s):
            return  # raw surrogate bytes are invalid UTF-8: one U+FFFD per bad byte, not per character
        assert d["node_names"] == want_names
        assert [c["percent"] for c in d["pod"]["containers"]] == [p or 0 for p in percents]
        assert [c["has"] for c in d["pod"]["containers"]] == [p is not None for p in percents]
        if all(not (0xD800 <= ord(c) < 0xE000) for k, v in annotations.items() for c in k + v):
            assert d["pod"]["annotations"] == annotations

    run()

    rng = random.Random(11)
    base = json.dumps({"Pod": x.pod_json([20, 40]), "NodeNames": ["node-%06d" % i for i in range(50)]}).encode()
    for _ in range(3000):
        b = bytearray(base)
        for _ in range(rng.randrange(1, 4)):
            i = rng.randrange(len(b))
            op = rng.randrange(3)
            if op == 0:
                b[i] = rng.randrange(256)
            elif op == 1:
                del b[i:i + rng.randrange(1, 20)]
            else:
                b[i:i] = bytes(rng.randrange(256) for _ in range(rng.randrange(1, 6)))
        d = x.decode_args(bytes(b))  # must come back (error text or not)
        assert isinstance(d["err"], str)
        try:
            ok = json.loads(bytes(b).decode("utf-8"))
        except Exception:
            ok = None
        if d["err"] == "" and ok is None:
            # the scanner accepted something Python rejects: only allowed for invalid UTF-8 (Go substitutes U+FFFD),
            # trailing bytes after the first value (json.Decoder reads one value) or lone-surrogate escapes
            try:
                json.loads(bytes(b).decode("utf-8", "replace"))
            except json.JSONDecodeError as e:
                assert "Extra data" in str(e), bytes(b)


def test_policy_file_and_go_durations():
    """getSyncPeriodFormPolicyConfig (controller.go:359-371) on the policy the reference deploys
    (deploy/dynamic-scheduler-node-annotator-cm.yaml:7-16, the text below is that ConfigMap's policy.yaml) and
    time.ParseDuration for the period strings"""
    from nano_gpu_scheduler_b200.routes import parse_duration, parse_policy

    deployed = """apiVersion: v1
kind: public-dynamic-scheduler
spec:
  syncPeriod:
    ##gpu usage
    - name: gpu_core_usage_avg
      period: 15s
    - name: gpu_memory_usage_avg
      period: 15s
"""
    assert parse_policy(deployed) == [("gpu_core_usage_avg", 15 * 10**9), ("gpu_memory_usage_avg", 15 * 10**9)]
    other = """
spec:
  priority:            # never read by the dealer (SURVEY a14)
    - name: gpu_core_usage_avg
      weight: 0.7
  syncPeriod:
  - name: "gpu_memory_usage_avg"   # quoted scalar, sequence at the parent's indent
    period: 1m30s
  - period: 500ms
    name: gpu_core_usage_avg
  - name: idle
status: {}
"""
    assert parse_policy(other) == [("gpu_memory_usage_avg", 90 * 10**9), ("gpu_core_usage_avg", 500 * 10**6), ("idle", 0)]
    assert parse_policy("kind: x\n") == [] and parse_policy("") == []
    with pytest.raises(ValueError):
        parse_policy("spec:\n  syncPeriod:\n    - name: a\n      period: soon\n")
    table = {"15s": 15 * 10**9, "1m30s": 90 * 10**9, "1.5h": 5400 * 10**9, "500ms": 5 * 10**8, "2us": 2000, "3µs": 3000, "7ns": 7, "0": 0,
             "-1m": -60 * 10**9, "+2s": 2 * 10**9, ".5s": 5 * 10**8, "1h2m3s4ms": 3723004 * 10**6}
    for s, want in table.items():
        assert parse_duration(s) == want, s
    for bad in ("", "15", "s", "1d", "1 s", "--1s", "1s2"):
        assert parse_duration(bad) is None, bad


def test_quantity_fuzz_against_fractions():
    """ParseQuantityValue against exact rational arithmetic: <digits>[.<digits>]<suffix>, rounded up, saturated at int64"""
    from fractions import Fraction
    import math

    rng = random.Random(23)
    dec = {"n": Fraction(1, 10**9), "u": Fraction(1, 10**6), "m": Fraction(1, 1000), "": Fraction(1), "k": Fraction(10**3), "M": Fraction(10**6),
           "G": Fraction(10**9), "T": Fraction(10**12), "P": Fraction(10**15), "E": Fraction(10**18)}
    bina = {"Ki": 2**10, "Mi": 2**20, "Gi": 2**30, "Ti": 2**40, "Pi": 2**50, "Ei": 2**60}
    for _ in range(4000):
        ip = str(rng.randrange(0, 10 ** rng.randrange(1, 8)))
        fp = "".join(rng.choice("0123456789") for _ in range(rng.randrange(0, 5)))
        num = ip + ("." + fp if fp or rng.random() < 0.1 else "")
        kind = rng.randrange(3)
        if kind == 0:
            suf = rng.choice(list(dec))
            mult = dec[suf]
        elif kind == 1:
            suf = rng.choice(list(bina))
            mult = Fraction(bina[suf])
        else:
            e = rng.randrange(-6, 7)
            suf = rng.choice("eE") + (rng.choice(["", "+"]) if e >= 0 else "") + str(e)
            mult = Fraction(10) ** e
        value = Fraction(num if not num.endswith(".") else num[:-1]) * mult
        want = min(math.ceil(value), 2**63 - 1)
        assert x.quantity(num + suf) == want, num + suf


def test_node_index_odd_names():
    """names of any length and content resolve to themselves (the route decoder hands over raw bytes)"""
    names = ["", "a", "a" * 7, "a" * 8, "a" * 9, "a" * 255, "node\x00with-nul", "ünï", "x" * 4096, "A", "a ", " a"]
    names += ["n%d" % i for i in range(5000)] + ["n%d." % i for i in range(5000)]
    body = json.dumps({"Pod": x.pod_json([10]), "NodeNames": names})
    d = x.decode_args(body)
    assert d["err"] == "" and d["node_names"] == names


def test_policy_parser_against_pyyaml():
    """generated policy files (entry order, indentation, comments, quoting, extra keys) read like PyYAML reads them"""
    import yaml

    from nano_gpu_scheduler_b200.routes import parse_duration, parse_policy

    rng = random.Random(5)
    units = ["ns", "us", "ms", "s", "m", "h"]
    for _ in range(300):
        ind = " " * rng.choice([2, 4])
        seq_ind = rng.choice(["", ind])  # a block sequence may sit at its parent's indent
        lines = ["apiVersion: v1", "kind: public-dynamic-scheduler  # trailing comment", "spec:"]
        if rng.random() < 0.5:
            lines += [ind + "priority:", ind + seq_ind + "- name: gpu_core_usage_avg", ind + seq_ind + "  weight: 0.5"]
        lines.append(ind + "syncPeriod:")
        want = []
        for _k in range(rng.randrange(0, 5)):
            name = rng.choice(["gpu_core_usage_avg", "gpu_memory_usage_avg", "x_%d" % rng.randrange(100)])
            dur = "".join("%d%s" % (rng.randrange(1, 90), u) for u in rng.sample(units, rng.randrange(1, 3)))
            q = rng.choice(['', '"', "'"])
            fields = ["name: %s%s%s" % (q, name, q), "period: %s" % dur]
            if rng.random() < 0.3:
                fields.append("note: ignored")
            rng.shuffle(fields)
            if rng.random() < 0.3:
                lines.append(ind + seq_ind + "# a comment between entries")
            lines.append(ind + seq_ind + "- " + fields[0])
            lines += [ind + seq_ind + "  " + f for f in fields[1:]]
            want.append((name, parse_duration(dur)))
        if rng.random() < 0.5:
            lines += ["status:", ind + "phase: x"]
        text = "\n".join(lines) + "\n"
        ref = yaml.safe_load(text)["spec"]["syncPeriod"] or []
        assert [(e["name"], parse_duration(str(e["period"]))) for e in ref] == want  # the generator and PyYAML agree
        assert parse_policy(text) == want, text
