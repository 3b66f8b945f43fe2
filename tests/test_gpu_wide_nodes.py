"""Nodes with 9..12 GPUs (the reference takes any count: pkg/utils/node.go:8-14, pkg/dealer/node.go:25-42; up to 12 cells
Go 1.16's sort.Sort is a gap-6 pass over (6,0)..(11,5) plus insertion sort).  The CUDA path keeps cells 8..11 in a second
plane; every verb and the single-pod scheduling call are compared with the oracle, bit for bit, on clusters that mix
all widths."""
import numpy as np
import pytest

import oracle as o
from tests.parity_util import NOW_S, load_ctx_pair

pytestmark = pytest.mark.gpu

POL = {"binpack": 0, "spread": 1, "random": 2}
WIDE = 12


@pytest.fixture(scope="module")
def ngs(built):
    import nano_gpu_scheduler_b200 as m

    return m


def mixed_pair(ngs, n, seed, p_used=0.6, with_rl=False, min_g=1):
    """the same cluster on the oracle and on the GPU: widths min_g..12, a few unknown rows, cells in steps of 5"""
    rng = np.random.default_rng(seed)
    oc = o.OracleCluster(n)
    d = ngs.GpuDealer(n, device=0)
    widths = rng.integers(min_g, WIDE + 1, n)
    widths[rng.random(n) < 0.02] = 0
    for i in range(n):
        G = int(widths[i])
        if G == 0:
            continue
        free = np.where(rng.random(G) < p_used, rng.integers(0, 21, G) * 5, 100).astype(np.int32)
        rl = rng.integers(-2, 3, G).astype(np.int8) if with_rl else None
        oc.node_set(i, free, rl)
        d.node_set_wide(i, G, free, rl)
    return oc, d, widths


def assert_same_nodes(oc, d, n):
    for i in range(n):
        G, fr, rl = d.node_get_wide(i)
        try:
            og, ofr, orl = oc.node_get(i)
        except IndexError:
            og, ofr, orl = 0, np.zeros(0), np.zeros(0)
        assert G == og, i
        assert fr[:G].tolist() == ofr.tolist(), (i, fr, ofr)
        assert rl[:G].tolist() == orl.tolist(), (i, rl, orl)


def demands(rng, trials):
    for t in range(trials):
        C = int(rng.integers(1, 13 if t % 3 == 0 else 6))
        yield (rng.integers(0, 13, C) * 5).astype(np.int32)


@pytest.mark.parametrize("policy", ["binpack", "spread", "random"])
@pytest.mark.parametrize("with_rl", [False, True])
def test_assume_with_plans_on_mixed_widths(ngs, policy, with_rl):
    n = 1500
    oc, d, _ = mixed_pair(ngs, n, 0x12A + with_rl, with_rl=with_rl)
    rng = np.random.default_rng(5)
    with d:
        for dem in demands(rng, 30):
            a = d.assume(dem, POL[policy], want_plan=True)
            b = oc.assume(dem, POL[policy], want_plan=True)
            for k in ("feasible", "reason", "score", "n_emitted", "plan"):
                assert np.array_equal(a[k], b[k]), (k, dem)
            # the streaming scan (all rows, no plans) and the candidate-list form answer alike
            a2 = d.assume(dem, POL[policy])
            idx = rng.permutation(n)[:400].astype(np.uint32)
            a3 = d.assume(dem, POL[policy], idx=idx)
            assert np.array_equal(a2["feasible"], b["feasible"]) and np.array_equal(a2["score"], b["score"])
            assert np.array_equal(a3["feasible"], b["feasible"][idx]) and np.array_equal(a3["score"], b["score"][idx])


def test_tie_order_of_twelve_equal_cells(ngs):
    """the gap-6 pass moves cells even when the slice is already sorted by key: the first pick among equal cells"""
    for G in range(7, WIDE + 1):
        for policy in ("binpack", "spread"):
            free = [100] * G
            oc = o.OracleCluster(1)
            oc.node_set(0, free)
            with ngs.GpuDealer(1, device=0) as d:
                d.node_set_wide(0, G, free)
                for dem in ([50], [50, 50, 50], [30, 20, 20, 10, 10, 10, 5]):
                    a = d.assume(dem, POL[policy], want_plan=True)
                    b = oc.assume(dem, POL[policy], want_plan=True)
                    assert a["plan"].tolist() == b["plan"].tolist(), (G, policy, dem)


@pytest.mark.parametrize("policy", ["binpack", "spread", "random"])
def test_filter_score_bind_cycles_on_mixed_widths(ngs, policy):
    """DealerImpl.Assume -> Score -> Bind per pod; the snapshots stay equal cell for cell, cells 8..11 included"""
    n = 300
    oc, d, _ = mixed_pair(ngs, n, 77, p_used=0.3)
    rng = np.random.default_rng(9)
    with d:
        for t, dem in enumerate(demands(rng, 150)):
            a = d.assume(dem, POL[policy])
            b = oc.assume(dem, POL[policy])
            assert np.array_equal(a["feasible"], b["feasible"]) and np.array_equal(a["score"], b["score"])
            sc = d.score(dem, POL[policy])
            assert np.array_equal(sc, b["score"])
            feas = np.flatnonzero(b["feasible"])
            if len(feas) == 0:
                continue
            best = feas[np.argmax(b["score"][feas])] if policy != "random" else feas[t % len(feas)]
            rc, plan, score = d.bind(int(best), dem, POL[policy])
            orc, oplan, oscore = oc.bind(int(best), dem, POL[policy])
            assert rc == orc == 0 and plan.tolist() == oplan.tolist() and score == oscore
        assert_same_nodes(oc, d, n)


def test_apply_release_and_release_batch_reach_cells_8_to_11(ngs):
    n = 40
    oc, d, widths = mixed_pair(ngs, n, 31, p_used=1.0, min_g=9)
    rng = np.random.default_rng(3)
    with d:
        pods = []
        for i in range(n):
            G = int(widths[i])
            if G == 0:
                continue
            gi = rng.integers(-1, G + 1, 4).astype(np.int32)  # -1: needs no GPU; G: one past the end
            pc = (rng.integers(0, 6, 4) * 5).astype(np.int32)
            pods.append((i, gi, pc))
        for i, gi, pc in pods:
            for sign in (+1, -1):
                a = d.node_apply(i, gi, pc, sign)
                b = oc.node_apply(i, gi, pc, sign)
                assert (a == 0) == (b == 0), (i, gi, pc, sign, a, b)
        assert_same_nodes(oc, d, n)
        # batched release
        node = np.array([p[0] for p in pods], dtype=np.uint32)
        gis = np.stack([np.clip(p[1], -1, None) for p in pods]).astype(np.int32)
        pcs = np.stack([p[2] for p in pods]).astype(np.int32)
        _, st = d.release_batch(node, gis, pcs, np.full(len(pods), 4, dtype=np.int32))
        for k, (i, _, _) in enumerate(pods):
            b = oc.node_apply(i, gis[k], pcs[k], -1)
            assert (st[k] == 0) == (b == 0)
        assert_same_nodes(oc, d, n)


@pytest.mark.parametrize("policy", ["binpack", "spread"])
def test_load_aware_on_mixed_widths(ngs, policy):
    """LoadUsage walks all twelve cards (allocate.go:173-195), RemainLoad of cells 8..11 is written back"""
    n = 800
    oc, d, widths = mixed_pair(ngs, n, 0x10AD)
    rng = np.random.default_rng(21)
    lg, lo = load_ctx_pair(ngs)
    with d:
        for i in range(n):
            if rng.random() < 0.1:
                continue  # no usage at all for this node
            for m in (0, 1):
                d.load_add_node(i, m)
                oc.load_add_node(i, m)
                for card in range(int(widths[i]) + 1):  # (one card more than the node has: ignored by LoadUsage)
                    if rng.random() < 0.15:
                        continue
                    v = float(rng.integers(0, 101)) / 100.0
                    ts = NOW_S - int(rng.integers(0, 400))
                    ok = rng.random() > 0.05
                    d.load_update(i, m, card, v, ts, value_ok=ok)
                    oc.load_update(i, m, card, v, ts, value_ok=ok)
        for dem in demands(rng, 12):
            a = d.assume(dem, POL[policy], load=lg, want_plan=True)
            b = oc.assume(dem, POL[policy], load=lo, want_plan=True)
            for k in ("feasible", "score", "n_emitted", "plan"):
                assert np.array_equal(a[k], b[k]), (k, dem)
            a2 = d.assume(dem, POL[policy], load=lg)
            assert np.array_equal(a2["score"], b["score"])
        assert_same_nodes(oc, d, n)
        # Bind after a load-aware filter: the plan comes from the RemainLoad that pass left behind
        dem = np.array([30, 20], dtype=np.int32)
        b = oc.assume(dem, POL[policy], load=lo)
        d.assume(dem, POL[policy], load=lg)
        for node in np.flatnonzero(b["feasible"])[:60]:
            ra = d.bind(int(node), dem, POL[policy], load=lg)
            rb = oc.bind(int(node), dem, POL[policy], load=lo)
            assert ra[0] == rb[0] and ra[1].tolist() == rb[1].tolist() and ra[2] == rb[2]
        assert_same_nodes(oc, d, n)


@pytest.mark.parametrize("policy", ["binpack", "spread", "random"])
def test_single_pod_scheduling_call_on_mixed_widths(ngs, policy):
    """ngs_schedule_batch with one pod (K7): scan, arg-max / seeded pick, bind on the device"""
    n = 700
    oc, d, _ = mixed_pair(ngs, n, 99, p_used=0.4, min_g=6)
    rng = np.random.default_rng(1)
    with d:
        for t, dem in enumerate(demands(rng, 120)):
            pc = np.zeros((1, 12), dtype=np.int32)
            pc[0, : len(dem)] = dem
            nc = np.array([len(dem)], dtype=np.int32)
            a = d.schedule_batch(pc, nc, [POL[policy]], seed=5, seq0=t)
            b = oc.schedule_batch(pc, nc, [POL[policy]], seed=5, seq0=t)
            assert a[0].tolist() == b[0].tolist() and a[1].tolist() == b[1].tolist(), (t, dem)
            assert a[2].tolist() == b[2].tolist(), (t, dem)
        assert_same_nodes(oc, d, n)


@pytest.mark.parametrize("policy", ["binpack", "spread", "random", "mixed"])
@pytest.mark.parametrize("n", [700, 40000])
def test_batches_on_mixed_widths(ngs, policy, n):
    """ngs_schedule_batch (K6): wide rows are filtered, planned and bound from the global tables, in order with the others"""
    oc, d, _ = mixed_pair(ngs, n, 1234 + n, p_used=0.5, min_g=5)
    rng = np.random.default_rng(8)
    P = 600
    pc = np.zeros((P, 6), dtype=np.int32)
    nc = rng.integers(1, 7, P).astype(np.int32)
    for p in range(P):
        pc[p, : nc[p]] = rng.integers(0, 13, nc[p]) * 5
    pol = rng.integers(0, 3, P).astype(np.int32) if policy == "mixed" else np.full(P, POL[policy], dtype=np.int32)
    with d:
        for lo in (0, 300):  # two calls: the second starts from what the first left in the tables
            a = d.schedule_batch(pc[lo:lo + 300], nc[lo:lo + 300], pol[lo:lo + 300], seed=3, seq0=lo)
            b = oc.schedule_batch(pc[lo:lo + 300], nc[lo:lo + 300], pol[lo:lo + 300], seed=3, seq0=lo)
            assert np.array_equal(a[0], b[0]), np.flatnonzero(a[0] != b[0])[:5]
            assert np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2])
        assert_same_nodes(oc, d, min(n, 3000))


@pytest.mark.parametrize("policy", ["binpack", "spread"])
def test_load_aware_batches_on_mixed_widths(ngs, policy):
    n = 1200
    oc, d, widths = mixed_pair(ngs, n, 4321, p_used=0.5, min_g=7)
    rng = np.random.default_rng(13)
    lg, lo_ = load_ctx_pair(ngs)
    with d:
        for i in range(n):
            for m in (0, 1):
                d.load_add_node(i, m)
                oc.load_add_node(i, m)
                for card in range(int(widths[i])):
                    if rng.random() < 0.3:
                        continue
                    v = float(rng.integers(0, 101)) / 100.0
                    ts = NOW_S - int(rng.integers(0, 400))
                    d.load_update(i, m, card, v, ts)
                    oc.load_update(i, m, card, v, ts)
        P = 200
        pc = (rng.integers(0, 9, (P, 3)) * 5).astype(np.int32)
        nc = np.full(P, 3, dtype=np.int32)
        pol = np.full(P, POL[policy], dtype=np.int32)
        a = d.schedule_batch(pc, nc, pol, load=lg)
        b = oc.schedule_batch(pc, nc, pol, load=lo_)
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2])
        assert_same_nodes(oc, d, n)


def test_limits(ngs):
    with ngs.GpuDealer(64, device=0) as d:
        with pytest.raises(ngs.NgsError) as e:
            d.node_upsert(0, 13)
        assert e.value.status == ngs.ERR_UNSUPPORTED
        gc = np.full(64, 8, dtype=np.uint8)
        d.snapshot_load(gc)
        gc[5] = 9
        with pytest.raises(ngs.NgsError):  # the bulk arrays are 8 cells wide
            d.snapshot_load(gc)
        d.node_upsert(7, 12)
        g, fr, rl = d.node_get_wide(7)
        assert g == 12 and fr.tolist() == [100] * 12 and rl.tolist() == [0] * 12
        d.node_upsert(7, 8)  # the node shrinks back
        g, fr, rl = d.node_get_wide(7)
        assert g == 8 and fr.tolist() == [100] * 8 + [0] * 4
