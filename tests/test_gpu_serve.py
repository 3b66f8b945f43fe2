"""K8, the resident single-pod kernel (ngs_serve_pod): the same decisions and table updates as the oracle (and hence as
ngs_schedule_batch), across idle exits, restarts and other calls in between."""
import time

import numpy as np
import pytest

import oracle as o
from tests.parity_util import NOW_S, fill_load_tables, load_ctx_pair, make_pair
from tests.test_gpu_wide_nodes import assert_same_nodes, mixed_pair

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(300)]


@pytest.fixture(scope="module")
def ngs(built):
    import nano_gpu_scheduler_b200 as m

    return m


def one(oc, dem, policy, load, seq):
    pc = np.zeros((1, 12), dtype=np.int32)
    pc[0, : len(dem)] = dem
    n, s, g = oc.schedule_batch(pc, np.array([len(dem)], dtype=np.int32), [policy], load=load, seed=9, seq0=seq)
    return int(n[0]), int(s[0]), g[0, : len(dem)].tolist()


def demands(rng, k):
    for _ in range(k):
        yield (rng.integers(0, 13, int(rng.integers(1, 6))) * 5).astype(np.int32)


@pytest.mark.parametrize("policy", [0, 1])
@pytest.mark.parametrize("load_on", [False, True])
def test_served_pods_match_the_oracle(ngs, policy, load_on):
    n = 30000
    oc, d, _, _ = make_pair(ngs, n, 0.5, 0x5E7 + policy)
    lg, lo = (None, None)
    rng = np.random.default_rng(3)
    with d:
        if load_on:
            fill_load_tables(oc, d, n, 0x5E7)
            lg, lo = load_ctx_pair(ngs)
        for seq, dem in enumerate(demands(rng, 300)):
            node, score, gidx = d.serve_pod(dem, policy, load=lg, seed=9, seq0=seq)
            assert (node, score, gidx.tolist()) == one(oc, dem, policy, lo, seq), (seq, dem)
        d.serve_stop()
        gc, fr, rl = d.snapshot_read()
        ogc, ofr, orl = oc.snapshot()
        assert np.array_equal(fr, ofr) and np.array_equal(rl, orl)


def test_idle_exit_restart_and_other_calls_in_between(ngs):
    n = 5000
    oc, d, _, _ = make_pair(ngs, n, 0.4, 77)
    rng = np.random.default_rng(5)
    with d:
        d.serve_idle_ms(2.0)
        launches0 = d.stats()["kernel_launches"]
        seq = 0
        for round_ in range(4):
            for dem in demands(rng, 25):
                node, score, gidx = d.serve_pod(dem, 0, seed=9, seq0=seq)
                assert (node, score, gidx.tolist()) == one(oc, dem, 0, None, seq)
                seq += 1
            if round_ == 0:
                time.sleep(0.05)  # the kernel leaves on its own; the next pod starts it again
            elif round_ == 1:
                dem = np.array([20, 20], dtype=np.int32)  # another verb in between stops it first
                a = d.assume(dem, 0)
                b = oc.assume(dem, 0)
                assert np.array_equal(a["feasible"], b["feasible"]) and np.array_equal(a["score"], b["score"])
            elif round_ == 2:
                d.node_apply(7, [0], [5], -1)
                oc.node_apply(7, [0], [5], -1)
        # random policy goes through the launched kernel from inside the call, same answer as the oracle
        for dem in demands(rng, 10):
            node, score, gidx = d.serve_pod(dem, 2, seed=9, seq0=seq)
            assert (node, score, gidx.tolist()) == one(oc, dem, 2, None, seq)
            seq += 1
        assert d.stats()["kernel_launches"] - launches0 < 60  # 100 served pods took a handful of launches, not 100
        d.serve_stop()
        d.serve_stop()  # idempotent
        gc, fr, rl = d.snapshot_read()
        assert np.array_equal(fr, oc.snapshot()[1])


def test_served_pods_on_clusters_with_wide_nodes(ngs):
    n = 3000
    oc, d, _ = mixed_pair(ngs, n, 4242, p_used=0.4, min_g=6)
    rng = np.random.default_rng(8)
    with d:
        for seq, dem in enumerate(demands(rng, 150)):
            policy = seq % 2
            node, score, gidx = d.serve_pod(dem, policy, seed=9, seq0=seq)
            assert (node, score, gidx.tolist()) == one(oc, dem, policy, None, seq), (seq, dem)
        d.serve_stop()
        assert_same_nodes(oc, d, n)


def test_a_pod_that_fits_nowhere_and_bad_arguments(ngs):
    with ngs.GpuDealer(64, device=0) as d:
        d.snapshot_load(np.full(64, 2, dtype=np.uint8))
        node, score, gidx = d.serve_pod([100, 100, 100], 0)
        assert node == -1 and score == 0 and gidx.tolist() == [-(2**31)] * 3
        with pytest.raises(ngs.NgsError):
            d.serve_pod([-5], 0)
        node, score, gidx = d.serve_pod([100, 100], 1)
        assert node == 0 and sorted(gidx.tolist()) == [0, 1]


@pytest.mark.parametrize("policy", [0, 1, 2])
@pytest.mark.parametrize("load_on", [False, True])
def test_verbs_through_the_resident_kernel_match_the_oracle(ngs, policy, load_on):
    """ngs_serve_verbs: Assume -> Score -> Bind per pod as requests to the resident kernel, every answer against the oracle"""
    n = 20000
    oc, d, _, _ = make_pair(ngs, n, 0.5, 0xBEE + policy, hetero=True)
    lg, lo = (None, None)
    rng = np.random.default_rng(17)
    with d:
        if load_on:
            fill_load_tables(oc, d, n, 0xBEE)
            lg, lo = load_ctx_pair(ngs)
        d.serve_verbs(True)
        launches0 = d.stats()["kernel_launches"]
        idx = rng.permutation(n)[:5000].astype(np.uint32)
        for t, dem in enumerate(demands(rng, 120)):
            use_idx = idx if t % 3 == 2 else None
            a = d.assume(dem, policy, idx=use_idx, load=lg)
            b = oc.assume(dem, policy, idx=use_idx, load=lo)
            for k in ("feasible", "reason", "score"):
                assert np.array_equal(a[k], b[k]), (t, k, dem)
            sc = d.score_view(dem, policy, idx=use_idx, load=lg) if t % 2 else d.score(dem, policy, idx=use_idx, load=lg)
            assert np.array_equal(sc, b["score"])
            feas = np.flatnonzero(b["feasible"])
            if len(feas) == 0:
                continue
            cand = feas if use_idx is None else use_idx[feas]
            best = int(cand[np.argmax(b["score"][feas])]) if policy != 2 else int(cand[t % len(cand)])
            ra = d.bind(best, dem, policy, load=lg)
            rb = oc.bind(best, dem, policy, load=lo)
            assert ra[0] == rb[0] == 0 and ra[1].tolist() == rb[1].tolist() and ra[2] == rb[2], (t, dem)
            if t == 40:  # a bind that cannot succeed and an unknown node: the same errors through the same path, nothing changes
                assert d.bind(best, [100] * 12, policy, load=lg)[0] == ngs.ERR_INFEASIBLE
                assert oc.bind(best, [100] * 12, policy, load=lo)[0] != 0
                assert d.bind(n + 5, dem, policy, load=lg)[0] == ngs.ERR_UNKNOWN_NODE
            if t % 20 == 5:  # releases and allocations in between are requests, too: the kernel keeps running
                assert d.node_apply(best, ra[1], dem, -1) == oc.node_apply(best, rb[1], dem, -1) == 0
                assert (d.node_apply(best, [0], [100], +1) == 0) == (oc.node_apply(best, [0], [100], +1) == 0)
                assert d.node_apply(n + 9, [0], [5], -1) == ngs.ERR_UNKNOWN_NODE
            if t % 10 == 7:  # ... and so are usage samples arriving from the metric sync
                node_u = int(rng.integers(0, n))
                for m in (0, 1):
                    if t % 20 == 7:
                        d.load_add_node(node_u, m)
                        oc.load_add_node(node_u, m)
                    for card in range(int(rng.integers(1, 9))):
                        v, ts = float(rng.integers(0, 101)) / 100.0, NOW_S - int(rng.integers(0, 300))
                        d.load_update(node_u, m, card, v, ts)
                        oc.load_update(node_u, m, card, v, ts)
        assert d.stats()["kernel_launches"] - launches0 < 12  # several hundred calls of six kinds, a handful of launches
        # a fresh score pass on its own (no filter call before it): also a request
        dem = np.array([15, 10], dtype=np.int32)
        assert np.array_equal(d.score(dem, policy, load=lg), oc.assume(dem, policy, load=lo)["score"])
        # plans travel through the resident kernel as well (packed, as from the launched scan): all of them, or the infeasible nodes' only
        launches1 = d.stats()["kernel_launches"]
        for dem in ([15, 10], [60, 55, 50], [100, 100, 95, 90]):
            a = d.assume(dem, policy, load=lg, want_plan=True)
            b = oc.assume(dem, policy, load=lo, want_plan=True)
            assert np.array_equal(a["plan"], b["plan"]) and np.array_equal(a["n_emitted"], b["n_emitted"])
            a = d.assume(dem, policy, load=lg, want_plan=True, failed_plans_only=True, idx=idx)
            b = oc.assume(dem, policy, load=lo, want_plan=True, idx=idx)
            bad = b["reason"] == ngs.REASON_INFEASIBLE
            assert np.array_equal(a["feasible"], b["feasible"]) and np.array_equal(a["score"], b["score"])
            assert np.array_equal(a["plan"][bad], b["plan"][bad]) and np.array_equal(a["n_emitted"][bad], b["n_emitted"][bad])
        assert d.stats()["kernel_launches"] - launches1 <= 2  # (the plan staging may have to grow once: that stops the kernel)
        d.serve_stop()
        gc, fr, rl = d.snapshot_read()
        ogc, ofr, orl = oc.snapshot()
        assert np.array_equal(fr, ofr) and np.array_equal(rl, orl)


def test_verbs_through_the_resident_kernel_on_wide_nodes(ngs):
    n = 2500
    oc, d, _ = mixed_pair(ngs, n, 777, p_used=0.4, min_g=4)
    rng = np.random.default_rng(23)
    with d:
        d.serve_verbs(True)
        for t, dem in enumerate(demands(rng, 80)):
            policy = t % 3
            a = d.assume(dem, policy)
            b = oc.assume(dem, policy)
            assert np.array_equal(a["feasible"], b["feasible"]) and np.array_equal(a["score"], b["score"])
            feas = np.flatnonzero(b["feasible"])
            if len(feas) == 0:
                continue
            best = int(feas[np.argmax(b["score"][feas])])
            ra = d.bind(best, dem, policy)
            rb = oc.bind(best, dem, policy)
            assert ra[0] == rb[0] == 0 and ra[1].tolist() == rb[1].tolist() and ra[2] == rb[2]
        d.serve_stop()
        assert_same_nodes(oc, d, n)


def test_requests_that_race_the_idle_exit(ngs):
    """an idle limit about as long as the gap between requests: the kernel keeps leaving just before, while or after a request
    arrives; every pod is still answered exactly once and exactly right"""
    n = 4000
    oc, d, _, _ = make_pair(ngs, n, 0.4, 1313)
    rng = np.random.default_rng(31)
    with d:
        d.serve_idle_ms(0.06)
        launches0 = d.stats()["kernel_launches"]
        for seq, dem in enumerate(demands(rng, 400)):
            if seq % 3:
                t_end = time.perf_counter() + float(rng.integers(0, 120)) * 1e-6
                while time.perf_counter() < t_end:
                    pass
            node, score, gidx = d.serve_pod(dem, seq % 2, seed=9, seq0=seq)
            assert (node, score, gidx.tolist()) == one(oc, dem, seq % 2, None, seq), (seq, dem)
        assert d.stats()["kernel_launches"] - launches0 > 20  # (it did leave and come back many times)
        d.serve_stop()
        assert np.array_equal(d.snapshot_read()[1], oc.snapshot()[1])
