"""GPU parity tests: the CUDA path, called through the C ABI, against the CPU oracle on the same
seeded inputs.  Everything here is integer / index work: the bar is bit-exact."""
import numpy as np
import pytest

import oracle as o
from tests.parity_util import (NOW_NS, NOW_S, assert_same_snapshot, fill_load_tables, load_ctx_pair, make_pair)

pytestmark = pytest.mark.gpu

POL = {"binpack": 0, "spread": 1, "random": 2}


@pytest.fixture(scope="module")
def ngs(built):
    import nano_gpu_scheduler_b200 as m

    return m


def single_node(ngs, free, rl=None):
    G = len(free)
    d = ngs.GpuDealer(1, device=0)
    row = np.zeros((1, 8), dtype=np.int32)
    row[0, :G] = free
    rlrow = None
    if rl is not None:
        rlrow = np.zeros((1, 8), dtype=np.int8)
        rlrow[0, :G] = rl
    d.snapshot_load(np.array([G], dtype=np.uint8), row, rlrow)
    return d


# ---- the reference's own known-answer tests, through the GPU ----------------------------------
@pytest.mark.parametrize("policy", ["binpack", "spread"])
def test_choose_kats_on_gpu(ngs, kats, policy):  # rater_test.go:133-401
    for c in kats[policy + "_choose"]["cases"]:
        with single_node(ngs, c["free"]) as d:
            r = d.assume(c["demand"], POL[policy], want_plan=True)
            assert bool(r["feasible"][0]) == (not c["expect_failed"])
            if not c["expect_failed"]:
                assert r["plan"][0].tolist() == c["expected"]
                assert r["reason"][0] == ngs.REASON_OK
            else:
                assert r["reason"][0] == ngs.REASON_INFEASIBLE and r["score"][0] == 0


def test_rate_kats_on_gpu(ngs, kats, derived):  # rater_test.go:9-131
    def score(policy, free):
        with single_node(ngs, free) as d:
            return int(d.assume([0], POL[policy])["score"][0])

    for c in kats["binpack_rate"]["cases"]:
        assert (score("binpack", c["gpus1"]) < score("binpack", c["gpus2"])) == c["first_lower"]
    for c in kats["spread_rate"]["cases"]:
        assert (score("spread", c["gpus1"]) > score("spread", c["gpus2"])) == c["first_preferred"]
    for v in derived["rates"]:
        assert score(v["policy"], v["free"]) == v["score"], v


def test_sample_choose_kats_on_gpu(ngs, kats):  # allocate_test.go:160-190
    for c in kats["sample_choose"]["cases"]:
        with single_node(ngs, c["free"]) as d:
            r = d.assume(c["demand"], POL["random"])
            assert bool(r["feasible"][0]) == c["success"]
            assert r["score"][0] == (100 if c["success"] else 0)


def test_gpu_resource_kats_on_gpu(ngs, kats):  # allocate_test.go:16-86
    g = kats["gpu_resource"]
    for c in g["sub"] + g["sub_if_available"]:
        with single_node(ngs, [c["origin"]]) as d:
            rc = d.node_apply(0, [0], [c["target"]], +1)
            assert (rc == 0) == (c["origin"] >= c["target"])
            assert d.snapshot_read()[1][0, 0] == c["expect"]
    for c in g["add"]:
        with single_node(ngs, [c["origin"]]) as d:
            assert d.node_apply(0, [0], [c["target"]], -1) == 0
            assert d.snapshot_read()[1][0, 0] == c["expect"]


def test_go116_tie_order_on_gpu(ngs, derived):
    """the GPU follows the pinned Go 1.16 sort.Sort (gap-6 pass), not the >= 1.19 rule"""
    for v in derived["sort_rule_choose"]:
        with single_node(ngs, v["free"]) as d:
            r = d.assume(v["demand"], POL[v["policy"]], want_plan=True)
            assert r["feasible"][0] == 1 and r["plan"][0].tolist() == v["go116"]


def test_fp64_truncation_points_on_gpu(ngs):
    """every used total 0..800 of an 8-GPU node: Binpack.Rate must truncate like Go's fp64"""
    n = 801
    free = np.full((n, 8), 100, dtype=np.int32)
    for used in range(n):
        left = used
        for g in range(8):
            take = min(100, left)
            free[used, g] -= take
            left -= take
    with ngs.GpuDealer(n, device=0) as d:
        d.snapshot_load(np.full(n, 8, dtype=np.uint8), free)
        sc = d.assume([0], POL["binpack"])["score"]
    want = np.array([int(float(u) / 800.0 * 100) - 8 for u in range(n)])
    assert np.array_equal(sc, want)
    assert sc[232] == 20 and sc[456] == 48 and sc[464] == 49


# ---- differential tests on random snapshots -----------------------------------------------------
@pytest.mark.parametrize("policy", ["binpack", "spread", "random"])
@pytest.mark.parametrize("hetero", [False, True])
def test_assume_random_snapshots(ngs, policy, hetero):
    n = 4096
    oc, d, _, _ = make_pair(ngs, n, 0.6, 0xA11CE + hetero, hetero=hetero)
    rng = np.random.default_rng(11)
    with d:
        for trial in range(24):
            C = int(rng.integers(1, 7))
            demand = (rng.integers(0, 13, C) * 5).astype(np.int32)
            if trial == 0:
                demand = np.array([0], dtype=np.int32)  # a pod that needs no GPU fits everywhere
            a = d.assume(demand, POL[policy], want_plan=True)
            b = oc.assume(demand, POL[policy], want_plan=True)
            for k in ("feasible", "reason", "score", "n_emitted", "plan"):
                assert np.array_equal(a[k], b[k]), (policy, trial, k, demand)
            # without the plan outputs the fast feasibility-only path runs: same answers
            a2 = d.assume(demand, POL[policy])
            assert np.array_equal(a2["feasible"], b["feasible"]) and np.array_equal(a2["score"], b["score"])


def test_assume_remain_load_in_sort_key(ngs):
    """non-load evaluation with a persisted RemainLoad: key = free + 50*rl (allocate.go:247)"""
    n = 2048
    rng = np.random.default_rng(5)
    free = (rng.integers(0, 11, (n, 8)) * 10).astype(np.int32)
    rl = rng.integers(0, 3, (n, 8)).astype(np.int8)
    oc = o.OracleCluster(n)
    for i in range(n):
        oc.node_set(i, free[i].astype(np.int64), rl[i].astype(np.int64))
    with ngs.GpuDealer(n, device=0) as d:
        d.snapshot_load(np.full(n, 8, dtype=np.uint8), free, rl)
        for trial in range(16):
            C = int(rng.integers(1, 5))
            demand = (rng.integers(1, 9, C) * 10).astype(np.int32)
            for pol in (0, 1):
                a = d.assume(demand, pol, want_plan=True)
                b = oc.assume(demand, pol, want_plan=True)
                for k in ("feasible", "score", "n_emitted", "plan"):
                    assert np.array_equal(a[k], b[k]), (pol, trial, k)


@pytest.mark.parametrize("policy", ["binpack", "spread"])
def test_assume_load_mode(ngs, policy):
    """isLoadSchedule=true: LoadUsage sums (fp64), RemainLoad write-back, stale / absent entries"""
    n = 3000
    oc, d, _, _ = make_pair(ngs, n, 0.5, 0x10AD)
    with d:
        fill_load_tables(oc, d, n, 0x10AD, stale_frac=0.1, absent_frac=0.05)
        gl, ol = load_ctx_pair(ngs)
        rng = np.random.default_rng(9)
        for trial in range(12):
            C = int(rng.integers(1, 5))
            demand = (rng.integers(0, 13, C) * 5).astype(np.int32)
            a = d.assume(demand, POL[policy], load=gl, want_plan=True)
            b = oc.assume(demand, POL[policy], load=ol, want_plan=True)
            for k in ("feasible", "score", "n_emitted", "plan"):
                assert np.array_equal(a[k], b[k]), (policy, trial, k)
            assert_same_snapshot(oc, d)  # RemainLoad side effect (allocate.go:193)
        # the window edge: one nanosecond decides (stats.go:48-53)
        ts = NOW_S - 100
        for i in range(8):
            for m in (0, 1):
                oc.load_update(i, m, 0, 0.95, ts)
                d.load_update(i, m, 0, 0.95, ts)
        for delta in (-1, 0, 1):
            win = (NOW_NS - ts * 10**9) + delta
            gl2, ol2 = load_ctx_pair(ngs, periods=[(0, win), (1, win)])
            a = d.assume([30], POL[policy], load=gl2, idx=np.arange(8, dtype=np.uint32))
            b = oc.assume([30], POL[policy], load=ol2, idx=np.arange(8, dtype=np.uint32))
            assert np.array_equal(a["score"], b["score"]) and np.array_equal(a["feasible"], b["feasible"])
            assert_same_snapshot(oc, d)


def test_load_usage_value_edge_cases(ngs):
    """values the reference skips (nodeusage.go:99-110) and the ceil(10u)/10 rounding"""
    vals = [0.0, 0.00001, 0.1, 0.10001, 0.3, 0.7, 0.70001, 0.99999, 1.0, 1.00001, -0.1, 0.55]
    n = len(vals)
    oc, d, _, _ = make_pair(ngs, n, 0.0, 1)
    with d:
        for i, v in enumerate(vals):
            for m in (0, 1):
                oc.load_add_node(i, m)
                for g in range(8):
                    vv = v if g % 2 == 0 else 0.2
                    oc.load_update(i, m, g, vv, NOW_S - 1, value_ok=(g != 3), ts_ok=(g != 5))
                    d.load_update(i, m, g, vv, NOW_S - 1, value_ok=(g != 3), ts_ok=(g != 5))
        gl, ol = load_ctx_pair(ngs)
        for pol in (0, 1):
            a = d.assume([10], pol, load=gl, want_plan=True)
            b = oc.assume([10], pol, load=ol, want_plan=True)
            assert np.array_equal(a["score"], b["score"]) and np.array_equal(a["plan"], b["plan"])
            assert_same_snapshot(oc, d)


def test_candidate_lists_and_unknown_nodes(ngs):
    n = 1000
    oc, d, gcount, free = make_pair(ngs, n, 0.5, 77)
    with d:
        # rows 10..19 are not known to the dealer
        gcount2 = gcount.copy()
        gcount2[10:20] = 0
        d.snapshot_load(gcount2, free)
        for i in range(10, 20):
            oc.node_upsert(i, 0)
        idx = np.array([5, 12, 999, 1000, 4_000_000_000, 5, 0, 19, 20], dtype=np.uint32)
        a = d.assume([30, 20], 0, idx=idx, want_plan=True)
        b = oc.assume([30, 20], 0, idx=idx, want_plan=True)
        for k in ("feasible", "reason", "score", "plan", "n_emitted"):
            assert np.array_equal(a[k], b[k]), k
        assert a["reason"][1] == ngs.REASON_UNKNOWN_NODE and a["reason"][3] == ngs.REASON_UNKNOWN_NODE
        # empty candidate list
        e = d.assume([30], 0, idx=np.zeros(0, dtype=np.uint32))
        assert len(e["feasible"]) == 0


def test_boundary_rejects_bad_input(ngs):
    with ngs.GpuDealer(8, device=0) as d:
        d.node_upsert(0, 8)
        with pytest.raises(ngs.NgsError) as e:
            d.assume([-5], 0)
        assert e.value.status == ngs.ERR_INVALID
        with pytest.raises(ngs.NgsError) as e:
            d.assume([10] * 13, 0)
        assert e.value.status == ngs.ERR_UNSUPPORTED
        with pytest.raises(ngs.NgsError) as e:
            d.node_upsert(1, 13)  # 9..12 are fine (tests/test_gpu_wide_nodes.py)
        assert e.value.status == ngs.ERR_UNSUPPORTED
        with pytest.raises(ngs.NgsError) as e:
            d.node_upsert(8, 8)
        assert e.value.status == ngs.ERR_UNKNOWN_NODE
        with pytest.raises(ngs.NgsError):
            d.assume([10], 7)
        # twelve containers (the insertion-sort regime limit) are fine
        r = d.assume([5] * 12, 0, want_plan=True)
        assert r["feasible"][0] == 1


def test_twelve_container_demand_sort(ngs):
    """demands of 7..12 containers exercise the gap-6 pass of the *demand* sort (rater.go:81)"""
    n = 512
    oc, d, _, _ = make_pair(ngs, n, 0.3, 0xD3)
    rng = np.random.default_rng(12)
    with d:
        for C in (7, 8, 9, 12):
            for _ in range(6):
                demand = (rng.integers(0, 4, C) * 10).astype(np.int32)
                for pol in (0, 1, 2):
                    a = d.assume(demand, pol, want_plan=True)
                    b = oc.assume(demand, pol, want_plan=True)
                    for k in ("feasible", "score", "plan", "n_emitted"):
                        assert np.array_equal(a[k], b[k]), (C, pol, k)


# ---- sequences: assume -> score -> bind -> release ---------------------------------------------
def test_bind_release_sequences(ngs):
    n = 300
    oc, d, _, _ = make_pair(ngs, n, 0.4, 0x5E0)
    rng = np.random.default_rng(21)
    live = []
    with d:
        for step in range(400):
            pol = int(rng.integers(0, 3))
            C = int(rng.integers(1, 4))
            demand = (rng.integers(0, 11, C) * 5).astype(np.int32)
            node = int(rng.integers(0, n))
            if rng.random() < 0.6 or not live:
                rc, gi, sc = d.bind(node, demand, pol)
                orc, ogi, osc = oc.bind(node, demand, pol)
                assert (rc == 0) == (orc == 0), step
                if rc == 0:
                    assert np.array_equal(gi, ogi) and sc == osc
                    live.append((node, demand, gi))
                else:
                    assert rc == ngs.ERR_INFEASIBLE
            else:
                node, demand, gi = live.pop(int(rng.integers(0, len(live))))
                assert d.node_apply(node, gi, demand, -1) == 0
                assert oc.node_apply(node, gi, demand, -1) == 0
            if step % 50 == 0:
                assert_same_snapshot(oc, d)
        assert_same_snapshot(oc, d)


def test_apply_error_paths(ngs):
    with single_node(ngs, [50, 20]) as d:
        oc = o.OracleCluster(1)
        oc.node_set(0, [50, 20])
        # second container does not fit: nothing is applied (all-or-nothing)
        assert d.node_apply(0, [0, 1], [30, 30], +1) == ngs.ERR_APPLY
        assert oc.node_apply(0, [0, 1], [30, 30], +1) != 0
        assert d.snapshot_read()[1][0, :2].tolist() == [50, 20] == oc.node_get(0)[1].tolist()
        # -1 entries are skipped (allocate.go:105-107)
        assert d.node_apply(0, [-1, 0], [70, 30], +1) == 0 and oc.node_apply(0, [-1, 0], [70, 30], +1) == 0
        # release with an index >= G: earlier containers stay released (allocate.go:120-131)
        assert d.node_apply(0, [1, 5], [10, 10], -1) == ngs.ERR_APPLY
        assert oc.node_apply(0, [1, 5], [10, 10], -1) != 0
        assert d.snapshot_read()[1][0, :2].tolist() == oc.node_get(0)[1].tolist() == [20, 30]
        # two containers on the same GPU are checked cumulatively
        assert d.node_apply(0, [1, 1], [20, 20], +1) == ngs.ERR_APPLY
        assert d.node_apply(0, [1, 1], [20, 10], +1) == 0
        assert d.node_apply(3, [0], [1], +1) == ngs.ERR_UNKNOWN_NODE


def test_score_uses_plan_cache(ngs):
    """NodeInfo.Score returns the Plan.Score cached by the filter pass (node.go:59-68) until the
    node changes; a different demand or a changed node is evaluated afresh."""
    n = 64
    oc, d, _, _ = make_pair(ngs, n, 0.5, 0xCAC4E)
    with d:
        fill_load_tables(oc, d, n, 3, stale_frac=0.0, absent_frac=0.0)
        gl, ol = load_ctx_pair(ngs)
        demand = [30, 20]
        a = d.assume(demand, 0, load=gl)
        # usage tables move on after the filter pass; the cached score must not
        for i in range(n):
            d.load_update(i, 0, 0, 1.0, NOW_S)
            d.load_update(i, 1, 0, 1.0, NOW_S)
        s_cached = d.score(demand, 0, load=gl)
        assert np.array_equal(s_cached, a["score"])
        # a node that changed is re-evaluated (cleanPlan, node.go:82,87,92)
        feas = np.flatnonzero(a["feasible"])
        target = int(feas[0])
        rc, gi, sc = d.bind(target, demand, 0, load=gl)
        assert rc == 0 and sc == a["score"][target]  # bind reuses the cached plan and score
        s2 = d.score(demand, 0, load=gl)
        same = np.ones(n, dtype=bool)
        same[target] = False
        assert np.array_equal(s2[same], a["score"][same])
        fresh = d.assume(demand, 0, load=gl, idx=np.array([target], dtype=np.uint32))
        assert s2[target] == fresh["score"][0]
        # a new demand opens a new cache generation: everything is computed fresh
        s3 = d.score([10], 0, load=gl)
        f3 = d.assume([10], 0, load=gl)
        assert np.array_equal(s3, f3["score"])


# ---- the batch kernel ---------------------------------------------------------------------------
def run_batch_pair(ngs, n, p_used, pods, c_rng, pct_rng, policy, seed, load=False, hetero=False, max_c=None):
    from nano_gpu_scheduler_b200 import workloads as wl

    oc, d, _, _ = make_pair(ngs, n, p_used, seed, hetero=hetero)
    pc, nc = wl.make_pods(pods, c_rng[0], c_rng[1], pct_rng[0], pct_rng[1], seed, max_c=max_c)
    with d:
        gl = ol = None
        if load:
            fill_load_tables(oc, d, n, seed)
            gl, ol = load_ctx_pair(ngs)
        node, score, gidx = d.schedule_batch(pc, nc, policy, load=gl, seed=seed, seq0=5)
        onode, oscore, ogidx = oc.schedule_batch(pc, nc, policy, load=ol, seed=seed, seq0=5)
        assert np.array_equal(node, onode)
        assert np.array_equal(score, oscore)
        assert np.array_equal(gidx, ogidx)
        assert_same_snapshot(oc, d)
        return node


def test_batch_config_c1(ngs):
    """BASELINE configs[0]: 10 nodes x 8 GPUs, 100 single-container pods, binpack"""
    node = run_batch_pair(ngs, 10, 0.0, 100, (1, 1), (1, 20), 0, 0xB2000001)
    assert (node >= 0).sum() > 50


@pytest.mark.parametrize("policy", [0, 1, 2])
def test_batch_small_clusters(ngs, policy):
    for n in (1, 7, 63, 64, 65, 1000, 9473):
        run_batch_pair(ngs, n, 0.5, 150, (1, 3), (0, 12), policy, 100 + n)


def test_batch_config_c2_full(ngs):
    """BASELINE configs[1] at its full size: 5k nodes, 10k pending pods, spread, single container (50 M oracle evaluations)"""
    run_batch_pair(ngs, 5000, 0.3, 10000, (1, 1), (1, 20), 1, 0xB2000002)


def test_batch_config_c3_scaled(ngs):
    """configs[2] shape: multi-container pods, binpack, load-aware"""
    run_batch_pair(ngs, 20000, 0.5, 400, (2, 4), (0, 12), 0, 0xB2000003, load=True, max_c=4)


def test_batch_hetero_gpu_counts(ngs):
    run_batch_pair(ngs, 3000, 0.4, 300, (1, 4), (0, 12), 0, 0x4E7, hetero=True)
    run_batch_pair(ngs, 3000, 0.4, 300, (1, 4), (0, 12), 1, 0x4E8, hetero=True, load=True)


def test_batch_mixed_policy_churn(ngs):
    """configs[4] shape: interleaved schedule / release with the policy cycling per pod"""
    from nano_gpu_scheduler_b200 import workloads as wl

    n = 5000
    oc, d, _, _ = make_pair(ngs, n, 0.5, 0xB2000005)
    rng = np.random.default_rng(55)
    live = []
    with d:
        for rnd in range(12):
            P = 200
            pc, nc = wl.make_pods(P, 1, 3, 1, 12, 0xB2000005 + rnd)
            pol = (np.arange(P) + rnd) % 3  # random / binpack / spread cycling
            node, score, gidx = d.schedule_batch(pc, nc, pol, seed=99, seq0=rnd * P)
            onode, oscore, ogidx = oc.schedule_batch(pc, nc, pol, seed=99, seq0=rnd * P)
            assert np.array_equal(node, onode) and np.array_equal(score, oscore) and np.array_equal(gidx, ogidx)
            for p in range(P):
                if node[p] >= 0:
                    live.append((int(node[p]), pc[p, : nc[p]].copy(), gidx[p, : nc[p]].copy()))
            # release a random third of the live pods (forget is bookkeeping only, dealer.go:311-319)
            for _ in range(len(live) // 3):
                nd, dem, gi = live.pop(int(rng.integers(0, len(live))))
                assert d.node_apply(nd, gi, dem, -1) == 0 and oc.node_apply(nd, gi, dem, -1) == 0
        assert_same_snapshot(oc, d)


def test_batch_matches_verb_by_verb(ngs):
    """one launch for P pods == P x (assume, score, arg-max, bind) through the single-pod verbs"""
    from nano_gpu_scheduler_b200 import workloads as wl

    n, P = 2000, 60
    gcount, free = wl.make_cluster(n, 0.5, 31)
    pc, nc = wl.make_pods(P, 1, 3, 0, 12, 31)
    with ngs.GpuDealer(n, device=0) as a, ngs.GpuDealer(n, device=0) as b:
        a.snapshot_load(gcount, free)
        b.snapshot_load(gcount, free)
        node, score, gidx = a.schedule_batch(pc, nc, 1)
        for p in range(P):
            dem = pc[p, : nc[p]]
            r = b.assume(dem, 1)
            s = b.score(dem, 1)
            cand = np.flatnonzero(r["feasible"])
            if len(cand) == 0:
                assert node[p] == -1
                continue
            best = cand[np.argmax(s[cand])]  # first maximum = lowest index
            assert node[p] == best and score[p] == s[best]
            rc, gi, sc = b.bind(int(best), dem, 1)
            assert rc == 0 and np.array_equal(gi, gidx[p, : nc[p]])
        assert np.array_equal(a.snapshot_read()[1], b.snapshot_read()[1])


def test_batch_rejects_bad_pods(ngs):
    with ngs.GpuDealer(100, device=0) as d:
        d.snapshot_load(np.full(100, 8, dtype=np.uint8))
        with pytest.raises(ngs.NgsError):
            d.schedule_batch(np.array([[-1, 0]], dtype=np.int32), [2], 0)
        with pytest.raises(ngs.NgsError):
            d.schedule_batch(np.array([[10, 0]], dtype=np.int32), [3], 0)
        node, _, _ = d.schedule_batch(np.zeros((0, 2), dtype=np.int32), np.zeros(0, dtype=np.int32), 0)
        assert len(node) == 0


def test_device_batch_with_one_bad_pod_schedules_the_rest(ngs):
    """device-resident batches are validated on the device: a malformed pod is reported as NGS_NODE_INVALID_POD and
    skipped, every other pod gets exactly the decision it would get in a batch without the bad one, the snapshot
    shows the good pods only, and the call's status says NGS_ERR_INVALID once"""
    import torch

    from nano_gpu_scheduler_b200 import workloads as wl

    n, P, max_c = 3000, 300, 4
    for bad_at, what in ((137, "percent"), (0, "ncont"), (299, "policy")):
        oc, d, _, _ = make_pair(ngs, n, 0.5, 0xBAD + bad_at)
        pc, nc = wl.make_pods(P, 1, 4, 0, 12, 0xBAD, max_c=max_c)
        pol = (np.arange(P) % 3).astype(np.int32)
        good = np.ones(P, dtype=bool)
        good[bad_at] = False
        bpc, bnc, bpol = pc.copy(), nc.copy(), pol.copy()
        if what == "percent":
            bpc[bad_at, 0] = -5
        elif what == "ncont":
            bnc[bad_at] = max_c + 1
        else:
            bpol[bad_at] = 7
        dev = torch.device("cuda", 0)
        with d:
            t_pc, t_nc, t_pol = (torch.from_numpy(a).to(dev) for a in (bpc, bnc, bpol))
            out = torch.full(((2 + max_c) * P,), 777, dtype=torch.int32, device=dev)
            b = out.data_ptr()
            d.schedule_batch_dev(P, t_pc.data_ptr(), t_nc.data_ptr(), max_c, t_pol.data_ptr(), b, b + 4 * P, b + 8 * P, seed=4, seq0=0)
            with pytest.raises(ngs.NgsError) as ei:
                d.sync()
            assert ei.value.status == ngs.ERR_INVALID
            d.sync()  # reported once
            o_ = out.cpu().numpy()
            node, score, gidx = o_[:P], o_[P:2 * P], o_[2 * P:].reshape(P, max_c)
            assert node[bad_at] == ngs._abi.NODE_INVALID_POD and score[bad_at] == 0 and (gidx[bad_at] == ngs.INT32_MIN).all()
            # the oracle schedules the good pods with the sequence numbers they had in the batch (the random policy's pick
            # is seeded per pod position): run it pod by pod
            for p_ in np.flatnonzero(good):
                on, os_, og = oc.schedule_batch(pc[p_:p_ + 1], nc[p_:p_ + 1], pol[p_:p_ + 1], seed=4, seq0=int(p_))
                assert node[p_] == on[0] and score[p_] == os_[0] and np.array_equal(gidx[p_], og[0]), (what, p_)
            assert_same_snapshot(oc, d)
            # a later batch on the same handle is not disturbed
            node2, _, _ = d.schedule_batch(pc[:50], nc[:50], pol[:50], seed=5)
            on2, _, _ = oc.schedule_batch(pc[:50], nc[:50], pol[:50], seed=5)
            assert np.array_equal(node2, on2)


def test_batch_chunking_across_launches(ngs):
    """more pods than one launch holds (4096): chunks continue the same sequence"""
    run_batch_pair(ngs, 400, 0.2, 5000, (1, 2), (1, 6), 2, 0xC4)
    run_batch_pair(ngs, 400, 0.2, 5000, (1, 2), (1, 6), 0, 0xC5)


# ---- full-size, size-independent properties ----------------------------------------------------
def test_full_size_c3_whole_bench_step(ngs):
    """100k nodes, load-aware binpack, 2..4 containers (BASELINE configs[2], the configuration the headline number is
    quoted on): ONE WHOLE bench step - 2048 pods in order - against the oracle, every decision and the snapshot bit for
    bit (205 M oracle evaluations, all host threads); then the size-independent properties on a longer run:
    (b) conservation, (c) every placed container fits where it was put, (d) schedule -> release restores the snapshot."""
    import os

    from nano_gpu_scheduler_b200 import workloads as wl

    n, P, K = 100_000, 3000, 2048
    gcount, free = wl.make_cluster(n, 0.5, 0xB2000003)
    pc, nc = wl.make_pods(P, 2, 4, 0, 12, 0xB2000003, max_c=4)
    with ngs.GpuDealer(n, device=0) as d:
        d.snapshot_load(gcount, free)
        tables = wl.make_load_tables(n, 0xB2000003, NOW_S)
        for m, (value, ts, absent) in enumerate(tables):
            ok = np.repeat(~absent[:, None], 8, axis=1).astype(np.uint8)
            d.load_update_bulk(m, value, ts, value_ok=ok, ts_ok=ok)
        gl, ol = load_ctx_pair(ngs)
        node, score, gidx = d.schedule_batch(pc[:K], nc[:K], 0, load=gl)
        # (a) the oracle on the whole step
        oc = o.OracleCluster(n)
        oc.load_snapshot(gcount, free)
        for m, (value, ts, absent) in enumerate(tables):
            oc.load_update_bulk(m, value, ts, ok=np.repeat(~absent[:, None], 8, axis=1).astype(np.uint8))
        onode, oscore, ogidx = oc.schedule_batch(pc[:K], nc[:K], 0, load=ol, n_threads=os.cpu_count() or 4)
        assert np.array_equal(node, onode) and np.array_equal(score, oscore) and np.array_equal(gidx, ogidx)
        assert_same_snapshot(oc, d)
        oc.close()
        # the rest of the run continues on the same snapshot
        node2, score2, gidx2 = d.schedule_batch(pc[K:], nc[K:], 0, load=gl, seq0=K)
        node, gidx = np.concatenate([node, node2]), np.concatenate([gidx, gidx2])
        _, after, _ = d.snapshot_read()
        placed = node >= 0
        assert placed.sum() > P // 2
        # (b) conservation
        removed = (free.astype(np.int64) - after).sum()
        assert removed == pc[placed].sum()
        # (c) replay the decisions on the host: every container fits at its turn
        cur = free.copy()
        for p in np.flatnonzero(placed):
            for k in range(nc[p]):
                g = gidx[p, k]
                if pc[p, k] == 0:
                    assert g == -1
                else:
                    assert 0 <= g < 8 and cur[node[p], g] >= pc[p, k]
                    cur[node[p], g] -= pc[p, k]
        assert np.array_equal(cur, after)
        # (d) round trip
        for p in np.flatnonzero(placed):
            assert d.node_apply(int(node[p]), gidx[p, : nc[p]], pc[p, : nc[p]], -1) == 0
        assert np.array_equal(d.snapshot_read()[1], free)


def test_full_size_c5_churn_slice(ngs):
    """BASELINE configs[4] at its full cluster size: 50k nodes, 6000 interleaved operations - schedule (policy cycling
    random / binpack / spread per pod, seeded replay) 60 %, release of a random live pod 35 %, forget 5 % (bookkeeping
    only, dealer.go:311-319) - every decision and the final snapshot against the oracle"""
    import os

    from nano_gpu_scheduler_b200 import workloads as wl

    n = 50_000
    oc, d, _, _ = make_pair(ngs, n, 0.5, 0xB2000005)
    rng = np.random.default_rng(0xC5)
    threads = os.cpu_count() or 4
    live, seq, ops = [], 0, 0
    with d:
        for rnd in range(10):
            P = 360  # 60 % of a 600-operation round
            pc, nc = wl.make_pods(P, 1, 3, 1, 12, 0xB2000005 + rnd, max_c=4)
            pol = ((np.arange(P) + seq) % 3 + 2) % 3  # random, binpack, spread, ...
            node, score, gidx = d.schedule_batch(pc, nc, pol, seed=0xB2000005, seq0=seq)
            onode, oscore, ogidx = oc.schedule_batch(pc, nc, pol, seed=0xB2000005, seq0=seq, n_threads=threads)
            assert np.array_equal(node, onode) and np.array_equal(score, oscore) and np.array_equal(gidx, ogidx), rnd
            seq += P
            for p in range(P):
                if node[p] >= 0:
                    live.append((int(node[p]), pc[p, : nc[p]].copy(), gidx[p, : nc[p]].copy()))
            k = min(210, len(live))  # 35 %: one batched Release call on the GPU, one by one on the oracle
            picks = sorted(rng.choice(len(live), size=k, replace=False).tolist(), reverse=True)
            rel = [live.pop(i) for i in picks]
            r_node = np.array([r[0] for r in rel], dtype=np.uint32)
            r_gi = np.full((k, 4), -1, dtype=np.int32)
            r_pc = np.zeros((k, 4), dtype=np.int32)
            r_nc = np.array([len(r[1]) for r in rel], dtype=np.int32)
            for i, r in enumerate(rel):
                r_pc[i, : len(r[1])] = r[1]
                r_gi[i, : len(r[2])] = r[2]
                assert oc.node_apply(r[0], r[2], r[1], -1) == 0
            rc, st = d.release_batch(r_node, r_gi, r_pc, r_nc)
            assert rc == 0 and not st.any()
            for _ in range(30):  # 5 %: Forget changes no capacity
                if live:
                    live.pop(int(rng.integers(0, len(live))))
            ops += P + k + 30
        assert ops >= 5000
        assert_same_snapshot(oc, d)


def test_full_size_million_nodes(ngs):
    """1M nodes (BASELINE configs[3], one shard): filter parity on a sampled candidate list and
    batch decisions against a numpy restatement of the single-container arg-max."""
    from nano_gpu_scheduler_b200 import workloads as wl

    n = 1_000_000
    gcount, free = wl.make_cluster(n, 0.5, 0xB2000004)
    with ngs.GpuDealer(n, device=0) as d:
        d.snapshot_load(gcount, free)
        r = d.assume([45], 0)
        # single container, RemainLoad 0: feasible iff some cell has >= 45 free; Binpack.Rate in fp64
        assert np.array_equal(r["feasible"].astype(bool), (free >= 45).any(axis=1))
        used = (100 - free.astype(np.int64)).sum(axis=1)
        want = (used.astype(np.float64) / 800.0 * 100.0).astype(np.int64) - 8
        assert np.array_equal(r["score"][r["feasible"] == 1], want[r["feasible"] == 1])
        pc, nc = wl.make_pods(64, 1, 1, 1, 20, 0xB2000004)
        node, score, gidx = d.schedule_batch(pc, nc, 0)
        # the oracle on the same 64 pods over all 1M nodes (and below an independent numpy restatement)
        import os

        oc = o.OracleCluster(n)
        oc.load_snapshot(gcount, free)
        onode, oscore, ogidx = oc.schedule_batch(pc, nc, 0, n_threads=os.cpu_count() or 4)
        assert np.array_equal(node, onode) and np.array_equal(score, oscore) and np.array_equal(gidx, ogidx)
        oc.close()
        cur = free.copy()
        for p in range(64):
            dem = pc[p, 0]
            feas = (cur >= dem).any(axis=1)
            u = (100 - cur.astype(np.int64)).sum(axis=1)
            sc = (u.astype(np.float64) / 800.0 * 100.0).astype(np.int64) - 8
            sc[~feas] = -(2**40)
            best = int(np.argmax(sc))
            assert node[p] == best and score[p] == sc[best]
            # binpack picks the fullest cell that still fits (ties are interchangeable here)
            assert cur[best, gidx[p, 0]] == cur[best][cur[best] >= dem].min()
            cur[best, gidx[p, 0]] -= dem
        assert np.array_equal(d.snapshot_read()[1], cur)


def test_assume_dev_outputs_stay_on_device(ngs):
    """ngs_assume_dev: same decisions as ngs_assume, outputs left in device memory (torch tensors)"""
    import torch

    n = 50_000
    oc, d, _, _ = make_pair(ngs, n, 0.5, 0xDE7)
    with d:
        fill_load_tables(oc, d, n, 0xDE7)
        gl, ol = load_ctx_pair(ngs)
        feas = torch.zeros(n, dtype=torch.uint8, device="cuda")
        score = torch.zeros(n, dtype=torch.int32, device="cuda")
        for pol, load_g, load_o, demand in ((0, None, None, [30]), (1, None, None, [40, 30, 20]), (0, gl, ol, [40, 30, 20])):
            d.assume_dev(demand, pol, feas.data_ptr(), score.data_ptr(), load=load_g)
            d.sync()
            want = oc.assume(demand, pol, load=load_o)
            assert np.array_equal(feas.cpu().numpy(), want["feasible"])
            got = score.cpu().numpy()
            assert np.array_equal(got[want["feasible"] == 1], want["score"][want["feasible"] == 1])
        # load scheduling off: the scan leaves the plan cache alone, so Score evaluates afresh and agrees
        d.assume_dev([30], 0, feas.data_ptr(), score.data_ptr())
        assert np.array_equal(d.score([30], 0), oc.assume([30], 0)["score"])


def test_two_handles_of_different_size_share_the_device(ngs):
    """a small handle must not shrink what a big one is allowed (the kernel's shared-memory limit is per device)"""
    from nano_gpu_scheduler_b200 import workloads as wl

    big_n, small_n = 400_000, 300
    gc_b, fr_b = wl.make_cluster(big_n, 0.5, 71)
    gc_s, fr_s = wl.make_cluster(small_n, 0.5, 72)
    pc, nc = wl.make_pods(40, 1, 3, 0, 12, 73)
    with ngs.GpuDealer(big_n, device=0) as big, ngs.GpuDealer(small_n, device=0) as small:
        big.snapshot_load(gc_b, fr_b)
        small.snapshot_load(gc_s, fr_s)
        n1, _, _ = big.schedule_batch(pc, nc, 0)
        oc = o.OracleCluster(small_n)
        oc.load_snapshot(gc_s, fr_s)
        n2, s2, g2 = small.schedule_batch(pc, nc, 1)
        on, os_, og = oc.schedule_batch(pc, nc, 1)
        assert np.array_equal(n2, on) and np.array_equal(s2, os_) and np.array_equal(g2, og)
        n3, _, _ = big.schedule_batch(pc, nc, 0)  # still launches with its large shared-memory footprint
        assert (n1 >= 0).all() and (n3 >= 0).all()


def test_release_batch_equals_one_by_one(ngs):
    """configs[4] shape at scale: schedule a mixed-policy stream, then give a random half back in ONE call"""
    from nano_gpu_scheduler_b200 import workloads as wl

    n, P = 20_000, 3000
    oc, d, _, free0 = make_pair(ngs, n, 0.5, 0xC5)
    pc, nc = wl.make_pods(P, 1, 3, 1, 12, 0xC5)
    pol = (np.arange(P) % 3).astype(np.int32)
    rng = np.random.default_rng(0xC5)
    with d:
        node, score, gidx = d.schedule_batch(pc, nc, pol, seed=5)
        onode, oscore, ogidx = oc.schedule_batch(pc, nc, pol, seed=5)
        assert np.array_equal(node, onode) and np.array_equal(gidx, ogidx)
        placed = np.flatnonzero(node >= 0)
        pick = rng.permutation(placed)[: len(placed) // 2]
        rc, st = d.release_batch(node[pick].astype(np.uint32), gidx[pick], pc[pick], nc[pick])
        assert rc == ngs.OK and (st == 0).all()
        for p in pick:
            assert oc.node_apply(int(node[p]), gidx[p, : nc[p]], pc[p, : nc[p]], -1) == 0
        assert_same_snapshot(oc, d)
        # per-pod error reporting: an index beyond the GPU count, an unknown node
        bad_node = np.array([3, 3, n + 7], dtype=np.uint32)
        bad_idx = np.array([[0, 9], [1, -1], [0, 0]], dtype=np.int32)
        bad_pc = np.array([[5, 5], [5, 0], [5, 5]], dtype=np.int32)
        rc, st = d.release_batch(bad_node, bad_idx, bad_pc, [2, 2, 2])
        assert rc != ngs.OK and st.tolist() == [ngs.ERR_APPLY, ngs.OK, ngs.ERR_UNKNOWN_NODE]
        assert oc.node_apply(3, [0, 9], [5, 5], -1) != 0 and oc.node_apply(3, [1, -1], [5, 0], -1) == 0
        assert_same_snapshot(oc, d)  # the first pod's container 0 stays released, as in the reference


def test_random_policy_seeded_replay(ngs):
    """the random policy (SURVEY a9'): same snapshot + same seed and sequence numbers -> the same picks; another
    seed -> other picks; every pick is a node the SampleRater plan search accepts (feasibility-set membership)"""
    from nano_gpu_scheduler_b200 import workloads as wl

    n, P = 4000, 300
    gcount, free = wl.make_cluster(n, 0.6, 0x5EED)
    pc, nc = wl.make_pods(P, 1, 3, 1, 14, 0x5EED)
    runs = []
    for seed in (1234, 1234, 99):
        with ngs.GpuDealer(n, device=0) as d:
            d.snapshot_load(gcount, free)
            runs.append(d.schedule_batch(pc, nc, ngs.POLICY_RANDOM, seed=seed, seq0=17))
    assert all(np.array_equal(a, b) for a, b in zip(runs[0], runs[1]))
    assert not np.array_equal(runs[0][0], runs[2][0])
    # membership: replay the first picks on the oracle and check each against the feasible set of its moment
    oc = o.OracleCluster(n)
    oc.load_snapshot(gcount, free)
    node, score, gidx = runs[0]
    for p in range(40):
        feas = oc.assume(pc[p, : nc[p]], o.POLICY_RANDOM)["feasible"]
        if node[p] < 0:
            assert feas.sum() == 0
            continue
        assert feas[node[p]] == 1 and score[p] == 100
        want = o.splitmix64(1234 + 17 + p) % int(feas.sum())
        assert node[p] == np.flatnonzero(feas)[want]
        rc, gi, _ = oc.bind(int(node[p]), pc[p, : nc[p]], o.POLICY_RANDOM)
        assert rc == 0 and np.array_equal(gi, gidx[p, : nc[p]])


def test_stats_counters(ngs):
    with ngs.GpuDealer(1000, device=0) as d:
        d.snapshot_load(np.full(1000, 8, dtype=np.uint8))
        s0 = d.stats()
        d.assume([10], 0)
        d.schedule_batch(np.full((5, 1), 10, dtype=np.int32), [1] * 5, 0)
        s1 = d.stats()
        assert s1["node_evals"] - s0["node_evals"] == 1000 + 5 * 1000
        assert s1["pods_scheduled"] - s0["pods_scheduled"] == 5
        assert s1["kernel_launches"] - s0["kernel_launches"] == 3
        assert s1["d2h_bytes"] > s0["d2h_bytes"] and s1["h2d_bytes"] > s0["h2d_bytes"]


@pytest.mark.parametrize("policy,load,hetero", [(0, False, False), (1, False, True), (0, True, True), (1, True, False), (2, False, True), (2, True, False)])
def test_single_pod_kernel_matches_oracle_and_batch_kernel(ngs, policy, load, hetero):
    """a one-pod call takes K7 (one ordinary launch, polled answer): pod by pod it must give what the oracle
    gives and what the batch kernel K6 gives for the same pods in one call, and leave the same snapshot"""
    from nano_gpu_scheduler_b200 import workloads as wl

    n, P, seed = 3000, 160, 0x0E + policy + 2 * load
    oc, d, gcount, free = make_pair(ngs, n, 0.55, seed, hetero=hetero)
    pc, nc = wl.make_pods(P, 1, 4, 0, 14, seed)
    with d, ngs.GpuDealer(n, device=0) as d6:
        d6.snapshot_load(gcount, free)
        gl = ol = None
        if load:
            fill_load_tables(oc, d, n, seed)
            fill_load_tables(o.OracleCluster(n), d6, n, seed)
            gl, ol = load_ctx_pair(ngs)
        node6, score6, gidx6 = d6.schedule_batch(pc, nc, policy, load=gl, seed=77, seq0=0)
        launches0 = d.stats()["kernel_launches"]
        for p in range(P):
            node, score, gidx = d.schedule_batch(pc[p : p + 1], nc[p : p + 1], policy, load=gl, seed=77, seq0=p)
            onode, oscore, ogidx = oc.schedule_batch(pc[p : p + 1], nc[p : p + 1], policy, load=ol, seed=77, seq0=p)
            assert node[0] == onode[0] == node6[p], p
            assert score[0] == oscore[0] == score6[p], p
            assert np.array_equal(gidx[0], ogidx[0]) and np.array_equal(gidx[0], gidx6[p]), p
        assert d.stats()["kernel_launches"] - launches0 == P  # one launch per pod: the K7 path was the one that ran
        assert (node6 >= 0).any()
        assert_same_snapshot(oc, d)
        assert_same_snapshot(oc, d6)


def test_single_pod_kernel_edge_cases(ngs):
    with ngs.GpuDealer(64, device=0) as d:
        d.snapshot_load(np.full(64, 2, dtype=np.uint8))
        # nothing fits: node -1, ScoreMin, no indexes; the exchange record must be clean for the next call
        node, score, gidx = d.schedule_batch(np.array([[100, 100, 100]], dtype=np.int32), [3], 0)
        assert node[0] == -1 and score[0] == 0 and (gidx[0] == np.iinfo(np.int32).min).all()
        # zero-percent containers get -1, unused slots INT32_MIN
        node, score, gidx = d.schedule_batch(np.array([[0, 30, 0, 0]], dtype=np.int32), [3], 0)
        assert node[0] == 0 and list(gidx[0]) == [-1, 0, -1, np.iinfo(np.int32).min]
        # a pod without containers binds to the best-rated node with an empty plan
        node, score, gidx = d.schedule_batch(np.zeros((1, 2), dtype=np.int32), [0], 0)
        assert node[0] >= 0 and (gidx[0] == np.iinfo(np.int32).min).all()
        # many calls in a row reuse the record; the random policy shares it
        for i in range(60):
            node, _, _ = d.schedule_batch(np.array([[100]], dtype=np.int32), [1], i % 3, seed=5, seq0=i)
        assert node[0] >= 0
        # random policy with nothing feasible, and on a cluster far larger than one row per thread
        node, score, gidx = d.schedule_batch(np.array([[100, 100, 100]], dtype=np.int32), [3], 2)
        assert node[0] == -1 and score[0] == 0
    n = 400_000
    from nano_gpu_scheduler_b200 import workloads as wl
    gcount, free = wl.make_cluster(n, 0.7, 31)
    oc = o.OracleCluster(n)
    oc.load_snapshot(gcount, free)
    with ngs.GpuDealer(n, device=0) as d:
        d.snapshot_load(gcount, free)
        for i in range(12):
            dem = np.array([[60, 45]], dtype=np.int32)
            got = d.schedule_batch(dem, [2], 2, seed=9, seq0=i)
            want = oc.schedule_batch(dem, [2], 2, seed=9, seq0=i)
            assert all(np.array_equal(a, b) for a, b in zip(got, want)), i
        assert_same_snapshot(oc, d)


def _overfull_cluster(n, seed):
    """free percents as stray Releases can leave them (GPUs.Release adds without an upper bound, allocate.go:120-131):
    most cells in 0..100, some in 101..300, a few up to NGS_MAX_FREE"""
    rng = np.random.default_rng(seed)
    gcount = rng.integers(1, 9, size=n).astype(np.uint8)
    kind = rng.random((n, 8))
    free = np.where(kind < 0.7, rng.integers(0, 101, (n, 8)), np.where(kind < 0.9, rng.integers(101, 301, (n, 8)), rng.integers(1000, 65536, (n, 8))))
    free = np.where(np.arange(8)[None, :] < gcount[:, None], free, 0).astype(np.int32)
    return gcount, free


DEMAND_VALUES = np.array([0, 5, 50, 100, 101, 150, 254, 255, 256, 300, 999, 40000, 65535], dtype=np.int32)


@pytest.mark.parametrize("policy", ["binpack", "spread", "random"])
def test_free_percent_beyond_100(ngs, policy):
    """rows above 100 % free: negative 'used' in Binpack.Rate (fp64 path, no table), large Spread rates, demands
    beyond one byte (the top-4 shortcut saturates at 255 and must fall back), u16 rows in the batch kernel"""
    n = 3000
    gcount, free = _overfull_cluster(n, 77)
    oc = o.OracleCluster(n)
    oc.load_snapshot(gcount, free)
    rng = np.random.default_rng(5)
    with ngs.GpuDealer(n, device=0) as d:
        d.snapshot_load(gcount, free)
        for trial in range(20):
            demand = rng.choice(DEMAND_VALUES, size=int(rng.integers(1, 5)))
            a = d.assume(demand, POL[policy], want_plan=True)
            b = oc.assume(demand, POL[policy], want_plan=True)
            for k in ("feasible", "reason", "score", "n_emitted", "plan"):
                assert np.array_equal(a[k], b[k]), (policy, trial, k, demand)
            a2 = d.assume(demand, POL[policy])  # the streaming scan
            assert np.array_equal(a2["feasible"], b["feasible"]) and np.array_equal(a2["score"], b["score"])
        P = 150
        pc = rng.choice(DEMAND_VALUES, size=(P, 4)).astype(np.int32)
        nc = rng.integers(1, 5, size=P).astype(np.int32)
        got = d.schedule_batch(pc, nc, POL[policy], seed=3)
        want = oc.schedule_batch(pc, nc, POL[policy], seed=3)
        assert all(np.array_equal(x, y) for x, y in zip(got, want))
        assert (got[0] >= 0).any()
        for p in range(40):  # and pod by pod through the single-pod kernel
            got = d.schedule_batch(pc[p : p + 1], nc[p : p + 1], POL[policy], seed=8, seq0=p)
            want = oc.schedule_batch(pc[p : p + 1], nc[p : p + 1], POL[policy], seed=8, seq0=p)
            assert all(np.array_equal(x, y) for x, y in zip(got, want)), p
        assert_same_snapshot(oc, d)


@pytest.mark.parametrize("policy", ["binpack", "spread", "random"])
def test_batch_with_up_to_twelve_containers(ngs, policy):
    """NGS_MAX_CONTAINERS-wide pods through K6 and K7 (Go's sort.Sort switches from insertion sort to quicksort
    above 12 elements, so 12 is the widest demand the reference sorts the simple way)"""
    from nano_gpu_scheduler_b200 import workloads as wl

    n, P = 1500, 120
    oc, d, _, _ = make_pair(ngs, n, 0.35, 0xC12, hetero=True)
    pc, nc = wl.make_pods(P, 6, 12, 0, 8, 0xC12, max_c=12)
    with d:
        got = d.schedule_batch(pc[:80], nc[:80], POL[policy], seed=1)
        want = oc.schedule_batch(pc[:80], nc[:80], POL[policy], seed=1)
        assert all(np.array_equal(x, y) for x, y in zip(got, want))
        for p in range(80, P):
            got = d.schedule_batch(pc[p : p + 1], nc[p : p + 1], POL[policy], seed=1, seq0=p)
            want = oc.schedule_batch(pc[p : p + 1], nc[p : p + 1], POL[policy], seed=1, seq0=p)
            assert all(np.array_equal(x, y) for x, y in zip(got, want)), p
        assert_same_snapshot(oc, d)


PERIOD_LISTS = [
    [(1, 320 * 10**9), (0, 320 * 10**9)],                                      # memory before core
    [(0, 320 * 10**9)],                                                        # core only
    [(1, 50 * 10**9)],                                                         # memory only, a window most samples miss
    [(0, 320 * 10**9), (0, 60 * 10**9), (1, 320 * 10**9), (1, 10**9)],          # NGS_MAX_PERIODS entries, metrics repeated
    [],                                                                        # load scheduling on, empty SyncPeriod
    [(0, 10**18), (1, -(10**12))],                                             # a window that always / never holds
]


@pytest.mark.parametrize("periods", PERIOD_LISTS, ids=lambda p: "periods%d" % len(p))
def test_load_mode_with_other_period_lists(ngs, periods):
    """PolicySpec.SyncPeriod lists other than the deployed (core, memory) pair take the general LoadUsage path
    (allocate.go:173-195 loops over whatever the list holds, in its order)"""
    from nano_gpu_scheduler_b200 import workloads as wl

    n = 2500
    oc, d, _, _ = make_pair(ngs, n, 0.5, 0x9E, hetero=True)
    with d:
        fill_load_tables(oc, d, n, 0x9E, stale_frac=0.15, absent_frac=0.05)
        gl, ol = load_ctx_pair(ngs, periods=periods)
        for policy in (0, 1):
            for demand in ([30], [20, 40, 0], [100, 100]):
                a = d.assume(demand, policy, load=gl, want_plan=True)
                b = oc.assume(demand, policy, load=ol, want_plan=True)
                for k in ("feasible", "score", "n_emitted", "plan"):
                    assert np.array_equal(a[k], b[k]), (policy, demand, k)
                a2 = d.assume(demand, policy, load=gl)
                assert np.array_equal(a2["score"], b["score"])
                assert_same_snapshot(oc, d)
        pc, nc = wl.make_pods(90, 1, 3, 1, 12, 0x9E)
        for policy in (0, 1, 2):
            got = d.schedule_batch(pc[:60], nc[:60], policy, load=gl, seed=2)
            want = oc.schedule_batch(pc[:60], nc[:60], policy, load=ol, seed=2)
            assert all(np.array_equal(x, y) for x, y in zip(got, want)), policy
            for p in range(60, 90):
                got = d.schedule_batch(pc[p : p + 1], nc[p : p + 1], policy, load=gl, seed=2, seq0=p)
                want = oc.schedule_batch(pc[p : p + 1], nc[p : p + 1], policy, load=ol, seed=2, seq0=p)
                assert all(np.array_equal(x, y) for x, y in zip(got, want)), (policy, p)
            assert_same_snapshot(oc, d)


@pytest.mark.parametrize("load", [False, True])
def test_many_rows_per_cta_all_policies(ngs, load):
    """clusters beyond 2048 rows per CTA run the batch kernel's all-warps-filter instantiation (k_schedule<true>):
    mixed policies, multi-container pods, zero-percent containers and load-aware scoring against the oracle"""
    from nano_gpu_scheduler_b200 import workloads as wl

    n, P = 330_000, 96  # 148 SMs -> 2230 rows per CTA
    oc, d, _, _ = make_pair(ngs, n, 0.6, 0xA77, hetero=True)
    pc, nc = wl.make_pods(P, 1, 4, 0, 14, 0xA77, max_c=4)
    pol = (np.arange(P) % 3).astype(np.int32)
    with d:
        gl = ol = None
        if load:
            fill_load_tables(oc, d, n, 0xA77)
            gl, ol = load_ctx_pair(ngs)
        got = d.schedule_batch(pc, nc, pol, load=gl, seed=11, seq0=3)
        want = oc.schedule_batch(pc, nc, pol, load=ol, seed=11, seq0=3, n_threads=8)
        for a, b, what in zip(got, want, ("node", "score", "gpu_idx")):
            assert np.array_equal(a, b), (what, np.flatnonzero((np.asarray(a) != np.asarray(b)).reshape(P, -1).any(axis=1))[:8])
        assert (got[0] >= 0).sum() > P // 2
        assert_same_snapshot(oc, d)


def test_score_view_is_score_without_the_copy(ngs):
    """ngs_score_view: the scores where the kernels left them — after a filter call (cache hit) and on its own (fresh pass)"""
    n = 5000
    oc, d, _, _ = make_pair(ngs, n, 0.6, 0x5C0)
    rng = np.random.default_rng(2)
    with d:
        for policy in (0, 1, 2):
            for trial in range(6):
                dem = (rng.integers(0, 13, int(rng.integers(1, 5))) * 5).astype(np.int32)
                want = oc.assume(dem, policy)["score"]
                if trial % 2 == 0:
                    d.assume(dem, policy)  # Score right after Assume: served from the filter call's scores
                v = d.score_view(dem, policy)
                assert not v.flags.writeable and np.array_equal(v, want)
                idx = rng.permutation(n)[:777].astype(np.uint32)
                assert np.array_equal(d.score_view(dem, policy, idx=idx), want[idx])
                assert np.array_equal(d.score(dem, policy), want)


@pytest.mark.parametrize("policy", [0, 1, 2])
def test_assume_with_plans_for_the_infeasible_only(ngs, policy):
    """ngs_assume_failed_plans: verdicts and scores as ngs_assume; plans / emitted counts for the infeasible candidates
    (the oracle's partial `indexes`), the rows of everyone else untouched"""
    n = 6000
    oc, d, _, _ = make_pair(ngs, n, 0.8, 0xFA11 + policy, hetero=True)
    rng = np.random.default_rng(4)
    with d:
        fill_load_tables(oc, d, n, 0xFA11)
        lg, lo = load_ctx_pair(ngs)
        for trial in range(12):
            dem = (rng.integers(2, 17, int(rng.integers(1, 6))) * 5).astype(np.int32)
            load_g, load_o = (lg, lo) if trial % 2 else (None, None)
            idx = rng.permutation(n)[:1500].astype(np.uint32) if trial % 3 == 0 else None
            a = d.assume(dem, policy, idx=idx, load=load_g, want_plan=True, failed_plans_only=True)
            b = oc.assume(dem, policy, idx=idx, load=load_o, want_plan=True)
            for k in ("feasible", "reason", "score"):
                assert np.array_equal(a[k], b[k]), (trial, k)
            bad = b["reason"] == ngs.REASON_INFEASIBLE
            assert bad.any() and (~bad).any()
            assert np.array_equal(a["n_emitted"][bad], b["n_emitted"][bad])
            assert np.array_equal(a["plan"][bad], b["plan"][bad])
            assert (a["n_emitted"][~bad] == 0).all() and (a["plan"][~bad] == -(2**31)).all()  # (as the wrapper initialised them)
