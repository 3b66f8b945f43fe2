"""The warp-cooperative building blocks of the batch kernel on the CPU: filter_pass + warp_top2, plan_group and
fresh_key_of_row from nano-gpu-scheduler_b200/csrc/ngs_k6_batch.cuh, compiled for the host (tests/host/k6_on_host.cpp)
and run on the lockstep warp simulator of tests/host/device_shims.h, against the oracle.  One simulated warp owns
all rows of a small "CTA".  Test harness only; k_schedule itself and its grid-wide exchange run in the -m gpu tests."""
import ctypes as C
import os
import shutil
import subprocess

import numpy as np
import pytest

import oracle as o

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "host", "k6_on_host.cpp")
LIB = os.path.join(HERE, "host", "libk6_on_host.so")
CSRC = os.path.join(HERE, "..", "nano-gpu-scheduler_b200", "csrc")
DEPS = [SRC, os.path.join(HERE, "host", "device_shims.h")] + [os.path.join(CSRC, f) for f in ("ngs_device.cuh", "ngs_rows.cuh", "ngs_k6_batch.cuh")]
INT32_MIN = -(2**31)


@pytest.fixture(scope="module")
def k6():
    cuda_inc = "/usr/local/cuda/include"
    if not os.path.isdir(cuda_inc):
        pytest.skip("no CUDA headers")
    if not os.path.exists(LIB) or max(os.path.getmtime(d) for d in DEPS) > os.path.getmtime(LIB):
        cxx = shutil.which("g++") or "g++"
        res = subprocess.run([cxx, "-O2", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off", "-I", cuda_inc, "-o", LIB, SRC],
                             capture_output=True, text=True)
        assert res.returncode == 0, res.stderr[-3000:]
    L = C.CDLL(LIB)
    L.k6h_error.restype = C.c_char_p
    L.k6h_key_node.argtypes = [C.c_uint64]
    L.k6h_key_node.restype = C.c_uint32
    L.k6h_key_score.argtypes = [C.c_uint64]
    return L


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def make_rows(rng, n, wide, with_rl):
    G = rng.integers(0, 9, n).astype(np.int32)  # 0 = unknown node
    vals = rng.choice([0, 5, 20, 50, 99, 100, 101, 254, 255, 256, 300, 4000, 65535], size=(n, 8)) if wide else rng.integers(0, 21, (n, 8)) * 5
    free = np.where(np.arange(8)[None, :] < G[:, None], vals, 0).astype(np.int32)
    rl = np.zeros((n, 8), dtype=np.int8)
    if with_rl:
        rl = np.where(np.arange(8)[None, :] < G[:, None], rng.integers(-2, 3, (n, 8)), 0).astype(np.int8)
    load_int = (rng.integers(0, 4, n) * (G > 0)).astype(np.int32) if with_rl else np.zeros(n, dtype=np.int32)
    return free, G, rl, load_int


def make_demand(rng, wide):
    if wide:
        return rng.choice([0, 5, 50, 100, 101, 255, 256, 300, 65535], size=int(rng.integers(0, 13))).astype(np.int32)
    return (rng.integers(0, 15, int(rng.integers(1, 7))) * 5).astype(np.int32)


def row_verdict(free, G, rl, load_int, r, demand, policy):
    """(feasible, score) of row r for the demand, from the oracle's bare functions"""
    g = int(G[r])
    if g == 0:
        return False, 0
    ok, _, _ = o.choose(policy, free[r, :g].astype(np.int64), demand, rl=rl[r, :g].astype(np.int64))
    if policy == o.POLICY_RANDOM:
        return ok, 100
    # rate() takes loadUsage as a float; int(loadUsage)/G is what the row keeps
    return ok, o.rate(policy, free[r, :g].astype(np.int64), float(int(load_int[r]) * g))


@pytest.mark.parametrize("policy", [o.POLICY_BINPACK, o.POLICY_SPREAD, o.POLICY_RANDOM])
@pytest.mark.parametrize("wide", [False, True])
def test_filter_pass_best_two_keys_and_feasibility_bits(k6, policy, wide):
    rng = np.random.default_rng(300 + 10 * policy + wide)
    for trial in range(250):
        n = int(rng.integers(1, 200))
        free, G, rl, load_int = make_rows(rng, n, wide, with_rl=bool(trial % 2))
        demand = make_demand(rng, wide)
        node0 = int(rng.integers(0, 1000)) * 1000
        m1, m2, cnt = C.c_uint64(0), C.c_uint64(0), C.c_uint32(0)
        bits = np.zeros((n + 31) // 32, dtype=np.uint32)
        rc = k6.k6h_filter(_p(free), _p(G), _p(rl), _p(load_int), n, node0, _p(demand) if len(demand) else None, len(demand), policy,
                           C.byref(m1), C.byref(m2), C.byref(cnt), _p(bits))
        assert rc == 0, k6.k6h_error()
        verdicts = [row_verdict(free, G, rl, load_int, r, demand, policy) for r in range(n)]
        if policy == o.POLICY_RANDOM:
            want_bits = [int(v[0]) for v in verdicts]
            got_bits = [(int(bits[r >> 5]) >> (r & 31)) & 1 for r in range(n)]
            assert got_bits == want_bits and cnt.value == sum(want_bits), (trial, n)
            continue
        # best two by (score desc, node asc)
        ranked = sorted(((-s, node0 + r) for r, (ok, s) in enumerate(verdicts) if ok))
        for k, m in enumerate((m1.value, m2.value)):
            if k < len(ranked):
                assert m != 0 and (-k6.k6h_key_score(m), k6.k6h_key_node(m)) == ranked[k], (trial, k, n, demand)
            else:
                assert m == 0, (trial, k)


def top4_word(cells):
    top = sorted((min(int(v), 255) for v in cells), reverse=True)[:4] + [0] * 4
    return top[0] | top[1] << 8 | top[2] << 16 | top[3] << 24


@pytest.mark.parametrize("policy", [o.POLICY_BINPACK, o.POLICY_SPREAD])
@pytest.mark.parametrize("wide", [False, True])
def test_plan_group_and_fresh_key(k6, policy, wide):
    """what control warp 0 works out for its candidate row while the exchange is in flight: the plan, the row as
    GPUs.Allocate leaves it, both rates and the top-4 word of that row, and its verdict for the NEXT pod"""
    rng = np.random.default_rng(900 + 10 * policy + wide)
    done = 0
    for trial in range(4000):
        n = 6
        free, G, rl, load_int = make_rows(rng, n, wide, with_rl=bool(trial % 2))
        G = np.maximum(G, 1)  # the candidate row exists
        free = np.where(np.arange(8)[None, :] < G[:, None], np.maximum(free, 0), 0).astype(np.int32)
        r = int(rng.integers(0, n))
        g = int(G[r])
        demand = make_demand(rng, wide)
        ok, idx, _ = o.choose(policy, free[r, :g].astype(np.int64), demand, rl=rl[r, :g].astype(np.int64))
        node = 5000 + r
        # fresh_key_of_row on the row as it stands must agree whether or not the plan fits
        has_next = bool(rng.integers(0, 2))
        nxt = make_demand(rng, wide)
        pol_next = int(rng.integers(0, 3))
        fr = np.zeros(8, dtype=np.int32)
        emit, fresh, key_now = C.c_uint64(0), C.c_uint64(0), C.c_uint64(0)
        n_emit, scB, scS, fresh_ok, ok_now, same = (C.c_int32(0) for _ in range(6))
        top4 = C.c_uint32(0)
        if not ok:
            continue  # plan_group is only ever asked about a row the filter found feasible
        rc = k6.k6h_plan(_p(free), _p(G), _p(rl), _p(load_int), n, r, node, _p(demand) if len(demand) else None, len(demand), policy,
                         _p(nxt) if len(nxt) else None, len(nxt) if has_next else -1, pol_next, _p(fr), C.byref(emit), C.byref(n_emit),
                         C.byref(top4), C.byref(scB), C.byref(scS), C.byref(fresh), C.byref(fresh_ok), C.byref(key_now), C.byref(ok_now),
                         C.byref(same))
        assert rc == 0, k6.k6h_error()
        assert same.value == 1  # every lane / group holds the same decision
        done += 1
        # the key of the row as it stands, for this pod
        want_ok, want_score = row_verdict(free, G, rl, load_int, r, demand, policy)
        assert bool(ok_now.value) == want_ok and k6.k6h_key_node(key_now.value) == node and k6.k6h_key_score(key_now.value) == want_score
        # the plan in container order, and the row after GPUs.Allocate
        got = np.full(max(len(demand), 1), INT32_MIN, dtype=np.int32)
        k6.k6h_unpermute(_p(demand) if len(demand) else None, len(demand), policy, C.c_uint64(emit.value), _p(got))
        assert got[: len(demand)].tolist() == idx and n_emit.value == len(demand), (trial, free[r, :g], rl[r, :g], demand)
        after = free[r].astype(np.int64).copy()
        for c, gi in enumerate(idx):
            if gi >= 0:
                after[gi] -= int(demand[c])
        assert fr[:g].tolist() == after[:g].tolist() and (fr[g:] == 0).all()
        # rates and top-4 of the row as it will stand
        lu = float(int(load_int[r]) * g)
        assert scB.value == o.rate(o.POLICY_BINPACK, after[:g], lu) and scS.value == o.rate(o.POLICY_SPREAD, after[:g], lu)
        assert top4.value == top4_word(after[:g])
        # ... and its verdict for the next pod on that future state
        if has_next:
            nok, _, _ = o.choose(pol_next, after[:g], nxt, rl=rl[r, :g].astype(np.int64))
            assert bool(fresh_ok.value) == nok, (trial, after[:g], nxt, pol_next)
            if nok and pol_next != o.POLICY_RANDOM:
                want = scS.value if pol_next == o.POLICY_SPREAD else scB.value
                assert k6.k6h_key_node(fresh.value) == node and k6.k6h_key_score(fresh.value) == want
            else:
                assert fresh.value == 0
        else:
            assert fresh.value == 0 and fresh_ok.value == 0
    assert done > 500


@pytest.mark.parametrize("policy", [o.POLICY_BINPACK, o.POLICY_SPREAD, o.POLICY_RANDOM])
@pytest.mark.parametrize("load", [False, True])
def test_row_evaluation_of_the_verb_kernels(k6, policy, load):
    """eval_row_values (ngs_rows.cuh) — what one thread of k_assume / k_scan_tma / K7 does with its row — against
    OracleCluster.assume on whole small clusters: feasibility, Plan.Score, plans and partial emission lists, and in
    load-aware mode the RemainLoad it writes back; both instantiations, with and without the usage-pair table"""
    import math

    from nano_gpu_scheduler_b200 import workloads as wl

    now_s = 1_790_000_000
    now_ns = now_s * 10**9 + 555_555_555
    rng = np.random.default_rng(40 + 3 * policy + load)
    n = 600
    gcount, free = wl.make_cluster(n, 0.55, 0xE7 + policy, hetero=True)
    gcount = gcount.copy()
    gcount[rng.integers(0, n, 12)] = 0  # a few unknown nodes
    free = np.where(np.arange(8)[None, :] < gcount[:, None], free, 0).astype(np.int32)
    tables = wl.make_load_tables(n, 0xE7, now_s, stale_frac=0.2, absent_frac=0.1)
    oc = o.OracleCluster(n)
    oc.load_snapshot(gcount, free)
    tenths = np.full((2, n, 8), 255, dtype=np.uint8)
    ts = np.zeros((2, n, 8), dtype=np.uint32)
    for m, (value, t, absent) in enumerate(tables):
        ok = np.repeat(~absent[:, None], 8, axis=1)
        oc.load_update_bulk(m, value, t, ok=ok.astype(np.uint8))
        good = ok & (value >= 0) & (value <= 1)
        tenths[m] = np.where(good, np.ceil(10 * value), 255).astype(np.uint8)
        ts[m] = t.astype(np.uint32)
    periods = wl.DEFAULT_PERIODS
    L = o.make_load_ctx(periods, now_ns) if load else None
    min_ts = np.array([max(0, math.floor((now_ns - w) / 10**9) + 1) for _, w in periods], dtype=np.uint32)
    G32 = gcount.astype(np.int32)
    rl = np.zeros((n, 8), dtype=np.int8)
    for trial in range(10):
        demand = (rng.integers(0, 13, int(rng.integers(1, 5))) * 5).astype(np.int32)
        want = oc.assume(demand, policy, load=L, want_plan=True)
        _, _, orl = oc.snapshot()
        for plan in (1, 0):
            for use_pair in (1, 0):
                feas = np.zeros(n, dtype=np.int32)
                score = np.zeros(n, dtype=np.int32)
                plans = np.full((n, len(demand)), INT32_MIN, dtype=np.int32)
                nem = np.zeros(n, dtype=np.int32)
                rl_run = rl.copy()
                k6.k6h_eval_rows(_p(free), _p(G32), _p(rl_run), _p(tenths), _p(ts), n, _p(demand), len(demand), policy, int(load), _p(min_ts),
                                 use_pair, plan, _p(feas), _p(score), _p(plans), _p(nem))
                assert np.array_equal(feas, want["feasible"]), (trial, plan, use_pair)
                assert np.array_equal(score, want["score"]), (trial, plan, use_pair)
                assert np.array_equal(rl_run, orl.astype(np.int8)), (trial, plan, use_pair)
                if plan:
                    assert np.array_equal(nem, want["n_emitted"]) and np.array_equal(plans, want["plan"]), (trial, use_pair)
        rl = orl.astype(np.int8)  # the next evaluation starts from what this one left behind
    oc.close()
