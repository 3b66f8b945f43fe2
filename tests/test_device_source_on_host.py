"""The kernels' own building blocks, compiled for the host, against the oracle — on the CPU.

tests/host/device_on_host.cpp includes nano-gpu-scheduler_b200/csrc/ngs_device.cuh (the source the CUDA kernels
include) and spells the few intrinsics it uses in plain C++, so the thread-serial logic — demand sort, packed-cell
plan search (choose_sorted), SampleRater, both Rate functions, the exact feasibility shortcuts, LoadUsage — can be
compared with the oracle on hundreds of thousands of rows without a GPU; choose_group8 (eight lanes per row) runs on
a lockstep warp simulator that also flags sync ops under divergent control flow.  This is a test harness only: the
product never loads it, and the kernels themselves are covered by the -m gpu tests."""
import ctypes as C
import os
import shutil
import subprocess

import numpy as np
import pytest

import oracle as o

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "host", "device_on_host.cpp")
LIB = os.path.join(HERE, "host", "libdevice_on_host.so")
CSRC = os.path.join(HERE, "..", "nano-gpu-scheduler_b200", "csrc")
HDRS = [os.path.join(CSRC, f) for f in ("ngs_device.cuh", "ngs_tables.cuh", "ngs_wide.cuh")]
INT32_MIN = -(2**31)
FEAS_NO, FEAS_YES, FEAS_UNDECIDED, FEAS_FALLBACK = 0, 1, 2, 3


@pytest.fixture(scope="module")
def doh():
    cuda_inc = "/usr/local/cuda/include"
    if not os.path.isdir(cuda_inc):
        pytest.skip("no CUDA headers")
    if not os.path.exists(LIB) or max(os.path.getmtime(f) for f in [SRC] + HDRS) > os.path.getmtime(LIB):
        cxx = shutil.which("g++") or "g++"
        # -ffp-contract=off: the fp64 sequence must stay add / mul / div in IEEE round-to-nearest, as on the device
        res = subprocess.run([cxx, "-O2", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off", "-I", cuda_inc, "-o", LIB, SRC],
                             capture_output=True, text=True)
        assert res.returncode == 0, res.stderr[-3000:]
    L = C.CDLL(LIB)
    L.doh_splitmix64.restype = C.c_uint64
    L.doh_splitmix64.argtypes = [C.c_uint64]
    L.doh_make_key.restype = C.c_uint64
    L.doh_make_key.argtypes = [C.c_int, C.c_uint32]
    L.doh_key_node.argtypes = [C.c_uint64]
    L.doh_key_node.restype = C.c_uint32
    L.doh_key_score.argtypes = [C.c_uint64]
    return L


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def eval_row(L, free, rl, demand, policy, load_int=0):
    G = len(free)
    f = np.zeros(8, dtype=np.int32)
    f[:G] = free
    r = np.zeros(8, dtype=np.int8)
    r[:G] = rl
    d = np.ascontiguousarray(demand, dtype=np.int32)
    out = np.zeros(6, dtype=np.int32)
    plan = np.full(max(len(d), 1), INT32_MIN, dtype=np.int32)
    L.doh_eval_row(_p(f), G, _p(r), _p(d), len(d), policy, load_int, _p(out), _p(plan))
    noplan = L.doh_feasible_noplan(_p(f), G, _p(r), _p(d), len(d), policy)
    return dict(ok=bool(out[0]), score=int(out[1]), ne=int(out[2]), cls_row=int(out[3]), cls_top4=int(out[4]), top4=int(out[5]) & 0xFFFFFFFF,
                plan=plan[: len(d)].tolist(), ok_noplan=bool(noplan))


def random_case(rng, wide):
    G = int(rng.integers(1, 9))
    if wide:
        free = rng.choice([0, 5, 20, 50, 99, 100, 101, 254, 255, 256, 300, 4000, 65535], size=G)
        demand = rng.choice([0, 5, 50, 100, 101, 255, 256, 300, 65535], size=int(rng.integers(0, 13)))
    else:
        free = rng.integers(0, 21, G) * 5  # many ties
        demand = rng.integers(0, 15, int(rng.integers(1, 7))) * 5
    rl = rng.integers(-2, 3, G) if rng.random() < 0.5 else np.zeros(G, dtype=np.int64)
    return free.astype(np.int64), rl.astype(np.int64), demand.astype(np.int32)


@pytest.mark.parametrize("policy", [o.POLICY_BINPACK, o.POLICY_SPREAD, o.POLICY_RANDOM])
@pytest.mark.parametrize("wide", [False, True])
def test_plan_search_rates_and_shortcuts_match_oracle(doh, policy, wide):
    rng = np.random.default_rng(1000 + 10 * policy + wide)
    undecided = decided = 0
    for trial in range(12000):
        free, rl, demand = random_case(rng, wide)
        got = eval_row(doh, free, rl, demand, policy)
        ok, idx, ne = o.choose(policy, free, demand, rl=rl)
        assert got["ok"] == ok and got["ok_noplan"] == ok, (trial, free, rl, demand)
        if ok:
            assert got["plan"] == idx, (trial, free, rl, demand)
        elif policy != o.POLICY_RANDOM:
            assert got["ne"] == ne and got["plan"][:ne] == idx[:ne], (trial, free, rl, demand)
        assert got["score"] == o.rate(policy, free), (trial, free)
        # the shortcuts may say "undecided", never the wrong thing
        if got["cls_row"] in (FEAS_NO, FEAS_YES):
            assert (got["cls_row"] == FEAS_YES) == ok, (trial, free, rl, demand, "classify_row")
            decided += 1
        else:
            assert got["cls_row"] == FEAS_UNDECIDED
            undecided += 1
        if policy != o.POLICY_RANDOM:
            # the four largest free values, saturated at 255, largest in the low byte
            top = sorted((min(int(v), 255) for v in free), reverse=True)[:4] + [0] * 4
            assert got["top4"] == top[0] | top[1] << 8 | top[2] << 16 | top[3] << 24
            if got["cls_top4"] in (FEAS_NO, FEAS_YES):
                assert (got["cls_top4"] == FEAS_YES) == ok, (trial, free, rl, demand, "classify_top4")
    assert decided > 0 and (policy == o.POLICY_RANDOM or undecided > 0)  # both outcomes of the shortcuts were exercised


def test_rates_with_load_term(doh):
    rng = np.random.default_rng(5)
    for _ in range(4000):
        G = int(rng.integers(1, 9))
        free = rng.integers(0, 101, G).astype(np.int64)
        load_usage = float(rng.integers(0, 170)) / 10.0
        load_int = int(load_usage) // G
        for policy in (o.POLICY_BINPACK, o.POLICY_SPREAD):
            got = eval_row(doh, free, np.zeros(G), [10], policy, load_int=load_int)
            assert got["score"] == o.rate(policy, free, load_usage), (free, load_usage, policy)


def test_demand_sort_matches_go_rule(doh):
    """sort.Sort of the demand (Go 1.16: gap-6 pass + insertion sort for <= 12 entries), processed largest first"""
    from tests import go_sort_py as gs

    rng = np.random.default_rng(8)
    for _ in range(4000):
        C_ = int(rng.integers(0, 13))
        pct = (rng.integers(0, 6, C_) * 20).astype(np.int32)  # many ties and zeros
        d = np.zeros(12, dtype=np.int32)
        order = np.zeros(12, dtype=np.int32)
        nz = C.c_int32(0)
        doh.doh_sort_demand(_p(pct) if C_ else None, C_, 0, _p(d), _p(order), C.byref(nz))
        cells = gs.Cells(pct.tolist(), [0] * C_, range(C_))
        gs.go116_sort(cells)
        want_order = [c[2] for c in cells.a][::-1]  # rater.go:82 walks from the largest down
        assert order[:C_].tolist() == want_order
        assert d[:C_].tolist() == [int(pct[i]) for i in want_order]
        assert nz.value == int((pct != 0).sum())


def _load_usage_py(tenths, ts, periods, G):
    """LoadUsage per card in SyncPeriod order, node sum in card order (allocate.go:173-195, rater.go:63-65), on the table encoding"""
    total, remain = 0.0, []
    for g in range(G):
        u = 0.0
        for metric, min_ts, never in periods:
            t10 = int(tenths[metric][g])
            if t10 == 255 or never or int(ts[metric][g]) < min_ts:
                continue
            u += min(t10, 10) / 10
        remain.append(2 - int(u))
        total += u
    return int(total) // G, remain


def test_load_usage_paths_agree_with_python(doh):
    rng = np.random.default_rng(21)
    for trial in range(6000):
        G = int(rng.integers(1, 9))
        tenths = rng.choice(np.arange(0, 11).tolist() + [255], size=(2, 8)).astype(np.uint8)
        ts = rng.integers(90, 111, size=(2, 8)).astype(np.uint32)
        if trial % 3 == 0:  # the deployed (core, memory) pair: all three code paths must agree
            periods = [(0, int(rng.integers(95, 106)), 0), (1, int(rng.integers(95, 106)), 0)]
            modes = (0, 1, 2)
        else:
            n = int(rng.integers(0, 5))
            periods = [(int(rng.integers(0, 2)), int(rng.integers(95, 106)), int(rng.random() < 0.15)) for _ in range(n)]
            modes = (0,)
        want_int, want_rl = _load_usage_py(tenths, ts, periods, G)
        metric = np.array([p[0] for p in periods] + [0] * 4, dtype=np.int32)
        min_ts = np.array([p[1] for p in periods] + [0] * 4, dtype=np.uint32)
        never = np.array([p[2] for p in periods] + [0] * 4, dtype=np.uint32)
        for mode in modes:
            rl_out = np.zeros(8, dtype=np.int32)
            got = doh.doh_load_usage(_p(tenths), _p(ts), len(periods), _p(metric), _p(min_ts), _p(never), G, mode, _p(rl_out))
            assert got == want_int and rl_out[:G].tolist() == want_rl, (trial, mode, G, periods)
            assert (rl_out[G:] == 0).all()


def test_keys_and_prng(doh):
    rng = np.random.default_rng(2)
    assert doh.doh_splitmix64(0) == o.splitmix64(0)
    for _ in range(2000):
        v = int(rng.integers(0, 2**63)) * 2 + int(rng.integers(0, 2))
        assert doh.doh_splitmix64(v) == o.splitmix64(v)
    # keys order by score first, then by LOWER node index; a feasible key is never 0
    keys = []
    for _ in range(3000):
        score, node = int(rng.integers(-20, 60000)), int(rng.integers(0, 2**32 - 1))
        k = doh.doh_make_key(score, node)
        assert k != 0 and doh.doh_key_node(k) == node and doh.doh_key_score(k) == score
        keys.append((k, score, node))
    keys.sort()
    for (k0, s0, n0), (k1, s1, n1) in zip(keys, keys[1:]):
        assert (s0, -n0) <= (s1, -n1)


def test_warp_simulator_reports_what_would_hang_a_warp(doh):
    """lanes that meet different sync ops (a ballot behind `ok &&`), or return while the others wait at one"""
    assert doh.doh_selftest_detects_hangs() == 3


@pytest.mark.parametrize("spread", [False, True])
@pytest.mark.parametrize("wide", [False, True])
def test_eight_lane_plan_search_matches_oracle(doh, spread, wide):
    """choose_group8 — the plan search of the batch kernel, one GPU cell per lane, four rows per warp — on a lockstep
    warp simulator: feasibility, emission order, the row as GPUs.Allocate leaves it, and no sync op under divergence"""
    doh.doh_warp_error.restype = C.c_char_p
    policy = o.POLICY_SPREAD if spread else o.POLICY_BINPACK
    rng = np.random.default_rng(77 + 2 * spread + wide)
    for trial in range(1500):
        if wide:
            demand = rng.choice([0, 5, 50, 100, 101, 255, 256, 300, 65535], size=int(rng.integers(0, 13))).astype(np.int32)
        else:
            demand = (rng.integers(0, 15, int(rng.integers(1, 7))) * 5).astype(np.int32)
        free = np.zeros((4, 8), dtype=np.int32)
        rl = np.zeros((4, 8), dtype=np.int8)
        G = rng.integers(0, 9, 4).astype(np.int32)  # G = 0: an unknown node in one of the four groups
        for r in range(4):
            vals = rng.choice([0, 5, 20, 50, 99, 100, 101, 254, 255, 256, 300, 4000, 65535], size=8) if wide else rng.integers(0, 21, 8) * 5
            free[r, : G[r]] = vals[: G[r]]
            if rng.random() < 0.5:
                rl[r, : G[r]] = rng.integers(-2, 3, G[r])
        C_ = len(demand)
        dptr = _p(demand) if C_ else None
        for plan in (1, 0):
            ok = np.zeros(4, dtype=np.int32)
            emit = np.zeros(4, dtype=np.uint64)
            ne = np.zeros(4, dtype=np.int32)
            fr = np.zeros((4, 8), dtype=np.int32)
            rc = doh.doh_choose_group8(_p(free), _p(G), _p(rl), dptr, C_, int(spread), plan, _p(ok), _p(emit), _p(ne), _p(fr))
            assert rc == 0, doh.doh_warp_error()
            for r in range(4):
                g = int(G[r])
                if g == 0:
                    # no cells: feasible only if nothing is demanded
                    assert bool(ok[r]) == (not (demand != 0).any()), (trial, r)
                    continue
                want_ok, idx, want_ne = o.choose(policy, free[r, :g].astype(np.int64), demand, rl=rl[r, :g].astype(np.int64))
                assert bool(ok[r]) == want_ok, (trial, r, free[r, :g], rl[r, :g], demand)
                if not plan or not want_ok:
                    continue
                got = np.full(max(C_, 1), INT32_MIN, dtype=np.int32)
                doh.doh_unpermute(dptr, C_, int(spread), C.c_uint64(int(emit[r])), _p(got))
                assert got[:C_].tolist() == idx and ne[r] == C_, (trial, r, free[r, :g], rl[r, :g], demand)
                after = free[r].astype(np.int64).copy()
                for c, gi in enumerate(idx):
                    if gi >= 0:
                        after[gi] -= int(demand[c])
                assert fr[r, :g].tolist() == after[:g].tolist(), (trial, r)


@pytest.mark.parametrize("wide", [False, True])
def test_eight_lane_search_with_one_demand_per_group_matches_oracle(doh, wide):
    """choose_group8_multi — feasibility of four (row, pod) pairs in one call: every 8-lane group has its own row, its own
    demand (different container counts, so the groups finish at different iterations) and its own policy, or nothing at all"""
    doh.doh_warp_error.restype = C.c_char_p
    rng = np.random.default_rng(4242 + wide)
    for trial in range(2500):
        free = np.zeros((4, 8), dtype=np.int32)
        rl = np.zeros((4, 8), dtype=np.int8)
        G = rng.integers(0, 9, 4).astype(np.int32)
        pct = np.zeros((4, 12), dtype=np.int32)
        Cn = np.zeros(4, dtype=np.int32)
        spread = rng.integers(0, 2, 4).astype(np.int32)
        for r in range(4):
            vals = rng.choice([0, 5, 20, 50, 99, 100, 101, 254, 255, 256, 300, 4000, 65535], size=8) if wide else rng.integers(0, 21, 8) * 5
            free[r, : G[r]] = vals[: G[r]]
            if rng.random() < 0.5:
                rl[r, : G[r]] = rng.integers(-2, 3, G[r])
            if rng.random() < 0.15:
                Cn[r] = -1  # this group asks nothing
                continue
            Cn[r] = int(rng.integers(0, 13 if wide else 7))
            if wide:
                pct[r, : Cn[r]] = rng.choice([0, 5, 50, 100, 101, 255, 256, 300, 65535], size=Cn[r])
            else:
                pct[r, : Cn[r]] = rng.integers(0, 15, Cn[r]) * 5
        ok = np.full(4, 7, dtype=np.int32)
        rc = doh.doh_choose_group8_multi(_p(free), _p(G), _p(rl), _p(pct), _p(Cn), _p(spread), _p(ok))
        assert rc == 0, doh.doh_warp_error()
        for r in range(4):
            if Cn[r] < 0:
                assert ok[r] == 0, (trial, r)
                continue
            demand = pct[r, : Cn[r]]
            g = int(G[r])
            if g == 0:
                assert bool(ok[r]) == (not (demand != 0).any()), (trial, r)
                continue
            policy = o.POLICY_SPREAD if spread[r] else o.POLICY_BINPACK
            want_ok, _, _ = o.choose(policy, free[r, :g].astype(np.int64), demand, rl=rl[r, :g].astype(np.int64))
            assert bool(ok[r]) == want_ok, (trial, r, free[r, :g], rl[r, :g], demand, spread[r])


@pytest.mark.parametrize("spread", [False, True])
def test_position_free_greedy_search_matches_oracle(doh, spread):
    """greedy_group8 — the search without positions: wherever it does not report an ambiguous tie, its verdict equals the
    oracle's, and (APPLY) the row it leaves equals, as a multiset of (free, RemainLoad) cells, the row the oracle's plan
    leaves; ties it reports must be real (two fitting cells of equal key and different free percent exist at some step)"""
    doh.doh_warp_error.restype = C.c_char_p
    policy = o.POLICY_SPREAD if spread else o.POLICY_BINPACK
    rng = np.random.default_rng(909 + spread)
    n_amb = n_tot = 0
    for trial in range(2500):
        demand = (rng.integers(0, 13, int(rng.integers(1, 6))) * 5).astype(np.int32)
        free = np.zeros((4, 8), dtype=np.int32)
        rl = np.zeros((4, 8), dtype=np.int8)
        G = rng.integers(0, 9, 4).astype(np.int32)
        for r in range(4):
            free[r, : G[r]] = rng.integers(0, 21, G[r]) * 5
            if rng.random() < 0.6:
                rl[r, : G[r]] = rng.integers(0, 3, G[r])
        for apply in (1, 0):
            st = np.zeros(4, dtype=np.int32)
            fr = np.zeros((4, 8), dtype=np.int32)
            rc = doh.doh_greedy_group8(_p(free), _p(G), _p(rl), _p(demand), len(demand), int(spread), apply, _p(st), _p(fr))
            assert rc == 0, doh.doh_warp_error()
            for r in range(4):
                g = int(G[r])
                n_tot += 1
                if st[r] & 2:
                    n_amb += 1  # (a tie can also arise from the placements themselves; the caller then runs the positional search)
                    assert (rl[r, :g] != rl[r, 0]).any(), (trial, r)  # ... but never on a row whose cells share one RemainLoad
                    continue
                if g == 0:
                    assert bool(st[r] & 1) == (not (demand != 0).any())
                    continue
                want_ok, idx, _ = o.choose(policy, free[r, :g].astype(np.int64), demand, rl=rl[r, :g].astype(np.int64))
                assert bool(st[r] & 1) == want_ok, (trial, r, free[r, :g], rl[r, :g], demand)
                if apply and want_ok:
                    after = free[r, :g].astype(np.int64).copy()
                    for c, gi in enumerate(idx):
                        if gi >= 0:
                            after[gi] -= int(demand[c])
                    got = sorted(zip(fr[r, :g].tolist(), rl[r, :g].tolist()))
                    want = sorted(zip(after.tolist(), rl[r, :g].tolist()))
                    assert got == want, (trial, r, free[r, :g], rl[r, :g], demand, idx)
    assert n_amb < 0.2 * n_tot  # the fast path has to be the common one


# ---- nodes with 9..12 GPUs: the one-thread evaluation of ngs_wide.cuh --------------------------------------------
def eval_wide(L, free, rl, demand, policy, usage=None, periods=(), do_rate=True):
    G = len(free)
    f = np.ascontiguousarray(free, dtype=np.int32)
    r = np.ascontiguousarray(rl, dtype=np.int8)
    d = np.ascontiguousarray(demand, dtype=np.int32)
    out = np.zeros(3, dtype=np.int32)
    plan = np.full(max(len(d), 1), INT32_MIN, dtype=np.int32)
    fa = np.zeros(12, dtype=np.int32)
    ra = np.zeros(12, dtype=np.int8)
    metric = np.array([p[0] for p in periods] + [0] * 8, dtype=np.int32)
    min_ts = np.array([p[1] for p in periods] + [0] * 8, dtype=np.uint32)
    never = np.array([p[2] for p in periods] + [0] * 8, dtype=np.uint32)
    tenths, ts = (None, None) if usage is None else usage
    L.doh_eval_wide_row(_p(f), G, _p(r), _p(d), len(d), policy, None if tenths is None else _p(tenths), None if ts is None else _p(ts),
                        len(periods), _p(metric), _p(min_ts), _p(never), int(do_rate), _p(out), _p(plan), _p(fa), _p(ra))
    return dict(ok=bool(out[0]), score=int(out[1]), ne=int(out[2]), plan=plan[: len(d)].tolist(), free_after=fa[:G].tolist(), rl_after=ra[:G].tolist())


@pytest.mark.parametrize("policy", [o.POLICY_BINPACK, o.POLICY_SPREAD, o.POLICY_RANDOM])
def test_wide_row_evaluation_matches_oracle(doh, policy):
    """G = 1..12 (the function is written for any width; the kernels send it rows of 9..12)"""
    rng = np.random.default_rng(4000 + policy)
    for trial in range(20000):
        G = int(rng.integers(1, 13)) if trial % 4 else int(rng.integers(9, 13))
        if trial % 5 == 0:
            free = rng.choice([0, 5, 20, 50, 99, 100, 101, 255, 300, 65535], size=G)
            demand = rng.choice([0, 5, 50, 100, 101, 255, 300, 65535], size=int(rng.integers(0, 13)))
        else:
            free = rng.integers(0, 21, G) * 5
            demand = rng.integers(0, 15, int(rng.integers(1, 13 if trial % 3 == 0 else 7))) * 5
        rl = rng.integers(-2, 3, G) if rng.random() < 0.5 else np.zeros(G, dtype=np.int64)
        got = eval_wide(doh, free, rl, demand, policy)
        ok, idx, ne = o.choose(policy, free.astype(np.int64), demand.astype(np.int32), rl=rl.astype(np.int64))
        assert got["ok"] == ok, (trial, free, rl, demand)
        if ok:
            assert got["plan"] == idx, (trial, free, rl, demand)
            after = free.astype(np.int64).copy()
            for k, gi in enumerate(idx):
                if gi >= 0:
                    after[gi] -= demand[k]
            assert got["free_after"] == after.tolist()
        elif policy != o.POLICY_RANDOM:
            assert got["ne"] == ne and got["plan"][:ne] == idx[:ne], (trial, free, rl, demand)
        assert got["score"] == o.rate(policy, free.astype(np.int64)), (trial, free)
        assert got["rl_after"] == rl.tolist()  # no load-aware pass: RemainLoad untouched


def test_wide_row_load_usage_matches_python_and_oracle_rate(doh):
    rng = np.random.default_rng(77)
    for trial in range(6000):
        G = int(rng.integers(1, 13))
        tenths = rng.choice(np.arange(0, 11).tolist() + [255], size=(2, 12)).astype(np.uint8)
        ts = rng.integers(90, 111, size=(2, 12)).astype(np.uint32)
        n = 2 if trial % 2 == 0 else int(rng.integers(0, 6))
        periods = [(int(rng.integers(0, 2)), int(rng.integers(95, 106)), int(rng.random() < 0.1)) for _ in range(n)]
        if trial % 2 == 0:
            periods = [(0, periods[0][1], 0), (1, periods[1][1], 0)]  # the deployed pair
        want_int, want_rl = _load_usage_py(tenths, ts, periods, G)
        free = (rng.integers(0, 21, G) * 5).astype(np.int64)
        total = 0.0
        for g in range(G):  # the fp64 node sum the oracle's Rate takes (rater.go:63-65)
            u = 0.0
            for metric, min_ts, never in periods:
                t10 = int(tenths[metric][g])
                if t10 == 255 or never or int(ts[metric][g]) < min_ts:
                    continue
                u += min(t10, 10) / 10
            total += u
        for policy in (o.POLICY_BINPACK, o.POLICY_SPREAD):
            got = eval_wide(doh, free, np.zeros(G), [10, 5], policy, usage=(tenths, ts), periods=periods)
            assert got["rl_after"] == want_rl, (trial, G, periods)
            assert got["score"] == o.rate(policy, free, total), (trial, free, total)
            # the plan uses the RemainLoad this very pass computed
            ok, idx, _ = o.choose(policy, free, np.array([10, 5], dtype=np.int32), rl=np.array(want_rl, dtype=np.int64))
            assert got["ok"] == ok and (not ok or got["plan"] == idx)
