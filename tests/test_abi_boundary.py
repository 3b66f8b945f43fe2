"""CPU-side checks of the drop-in boundary: the CUDA library builds, loads, and exports every
symbol include/ngs_dealer.h declares; host-side helpers behave like the Go calls they replace.
No compute calls here (no GPU)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "ngs_dealer.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ngs_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(built):
    import nano_gpu_scheduler_b200 as ngs

    lib = ngs._abi.load()
    syms = header_symbols()
    assert len(syms) >= 20
    for s in syms:
        assert hasattr(lib, s), s
    assert sorted(ngs._abi.SYMBOLS) == syms
    assert lib.ngs_abi_version() == 3


def test_no_torch_or_oracle_in_the_product_library(built):
    """the C ABI carries no torch types and the product never links the oracle"""
    import subprocess

    import nano_gpu_scheduler_b200 as ngs

    out = subprocess.run(["ldd", ngs._abi.LIB_PATH], capture_output=True, text=True).stdout
    assert "torch" not in out and "dealer_oracle" not in out
    for root, _dirs, files in os.walk(os.path.join(ROOT, "nano-gpu-scheduler_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                txt = open(os.path.join(root, f)).read()
                assert "import oracle" not in txt and "dealer_oracle" not in txt, f


def test_status_strings(built):
    import nano_gpu_scheduler_b200 as ngs

    lib = ngs._abi.load()
    assert lib.ngs_strerror(0) == b"ok"
    assert b"CUDA" in lib.ngs_strerror(ngs.ERR_CUDA)
    assert lib.ngs_strerror(12345) == b"unknown status"


def test_create_fails_loudly_without_gpu(built):
    import torch

    import nano_gpu_scheduler_b200 as ngs

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(ngs.NgsError) as e:
        ngs.GpuDealer(16)
    assert e.value.status == ngs.ERR_CUDA


def test_parse_usage_like_strconv_parsefloat(built):  # nodeusage.go:103
    import nano_gpu_scheduler_b200 as ngs

    assert ngs.parse_usage("0.70000") == (True, 0.7)
    assert ngs.parse_usage("1") == (True, 1.0)
    assert ngs.parse_usage("1e-1") == (True, 0.1)
    for bad in ["", " 0.5", "0.5 ", "abc", "0.5x", "1e400"]:
        assert ngs.parse_usage(bad)[0] is False, bad


def test_parse_update_time_like_go_layout(built):  # stats.go:30-44, type.go:12
    import calendar

    import nano_gpu_scheduler_b200 as ngs

    ok, ts = ngs.parse_update_time("2024-03-01T12:00:00Z")
    # rendered in Asia/Shanghai (UTC+8): 12:00 there is 04:00 UTC
    assert ok and ts == calendar.timegm((2024, 3, 1, 4, 0, 0))
    assert ngs.parse_update_time("2024-02-29T23:59:59Z")[0]
    assert ngs.parse_update_time("2024-03-01T12:00:00.250Z")[0]
    for bad in ["", "abcd", "2024-03-01 12:00:00Z", "2024-03-01T12:00:00", "2023-02-29T00:00:00Z",
                "2024-13-01T00:00:00Z", "2024-03-01T24:00:00Z", "2024-3-1T12:00:00Z"]:
        assert ngs.parse_update_time(bad)[0] is False, bad


def test_load_ctx_layout_matches_header(built):
    import nano_gpu_scheduler_b200 as ngs

    assert C.sizeof(ngs._abi.LoadCtx) == 4 + 4 * 8 + 4 + 8 * 8 + 8  # int32, int32[8], pad, int64[8], int64
    assert C.sizeof(ngs._abi.Stats) == 40


def test_parse_fuzz_against_python(built):
    """random well-formed update times (any second from 1992 to 2100, the years Asia/Shanghai is plainly UTC+8) and
    usage strings as the metric sync writes them ('%.5f'), plus damaged variants of both"""
    import calendar
    import random

    import nano_gpu_scheduler_b200 as ngs

    rng = random.Random(99)
    lo, hi = calendar.timegm((1992, 1, 1, 0, 0, 0)), calendar.timegm((2100, 1, 1, 0, 0, 0))
    for _ in range(3000):
        t = rng.randrange(lo, hi)
        y, mo, d, h, mi, s = __import__("time").gmtime(t + 8 * 3600)[:6]
        text = "%04d-%02d-%02dT%02d:%02d:%02dZ" % (y, mo, d, h, mi, s)
        assert ngs.parse_update_time(text) == (True, t), text
        i = rng.randrange(len(text))
        broken = text[:i] + rng.choice("x :-/T") + text[i + 1:]
        if broken != text:
            ok, _ = ngs.parse_update_time(broken)
            assert ok is False, broken
    for _ in range(3000):
        v = rng.random() * rng.choice([1.0, 1.0, 2.0])
        text = "%.5f" % v
        assert ngs.parse_usage(text) == (True, float(text)), text
        text2 = repr(rng.uniform(-3, 3))
        assert ngs.parse_usage(text2) == (True, float(text2)), text2
