"""In-tree build of the CUDA library (nvcc, sm_100a only).  Used by __graft_entry__.build()."""
from __future__ import annotations

import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libngs_dealer.so")
SOURCES = ["ngs_dealer.cu", "ngs_parse.cpp"]
HEADERS = [os.path.join(CSRC, f) for f in ("ngs_device.cuh", "ngs_tables.cuh", "ngs_wide.cuh", "ngs_rows.cuh", "ngs_k1_scan.cuh", "ngs_k3_rows.cuh", "ngs_k6_batch.cuh", "ngs_k7_one.cuh", "ngs_k8_serve.cuh")] + [
    os.path.join(HERE, "..", "include", "ngs_dealer.h")]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-shared", "-Xcompiler", "-fPIC",
]


def find_nvcc() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: the dealer has no CPU fallback and cannot be built without CUDA")


def is_stale() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, s) for s in SOURCES] + HEADERS
    return any(os.path.getmtime(d) > t for d in deps)


HOST_DIR = os.path.join(HERE, "host")
HOST_LIB = os.path.join(HERE, "libngs_dealer_host.so")
HOST_TEST = os.path.join(HERE, "host", "test_dealer_host")


def build_host(force: bool = False) -> str:
    """C++ mirror of the reference's Dealer interface (host/dealer_host.cpp), the extender wire codec
    (host/extender_codec.cpp) and their test program"""
    srcs = [os.path.join(HOST_DIR, f) for f in ("dealer_host.cpp", "extender_codec.cpp", "extender_codec_c.cpp")]
    hdrs = [os.path.join(HOST_DIR, f) for f in ("dealer_host.h", "extender_codec.h")]
    test_src = os.path.join(HERE, "..", "tests", "host", "test_dealer_host.cpp")
    deps = [*srcs, *hdrs, test_src, LIB]
    if force or not os.path.exists(HOST_TEST) or any(os.path.getmtime(d) > os.path.getmtime(HOST_TEST) for d in deps):
        cxx = shutil.which("g++") or "g++"
        common = ["-O2", "-std=c++17", "-fPIC", "-Wall", "-pthread", "-I", os.path.join(HERE, "..", "include")]
        res = subprocess.run([cxx, *common, "-shared", "-o", HOST_LIB, *srcs, "-L", HERE, "-lngs_dealer", "-Wl,-rpath,$ORIGIN"],
                             capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError("g++ (host mirror) failed:\n" + res.stdout + res.stderr)
        res = subprocess.run([cxx, *common, "-o", HOST_TEST, test_src, *srcs[:2], "-L", HERE, "-lngs_dealer", "-Wl,-rpath,$ORIGIN/.."],
                             capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError("g++ (host test) failed:\n" + res.stdout + res.stderr)
    return HOST_TEST


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not is_stale():
        return LIB
    cmd = [find_nvcc(), *NVCC_FLAGS, "-o", LIB] + [os.path.join(CSRC, s) for s in SOURCES]
    if verbose:
        cmd.insert(1, "-Xptxas")
        cmd.insert(2, "-v")
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
    if verbose:
        print(res.stderr)
    return LIB
