"""The oracle against every known-answer vector the reference's own tests hold for the path
(tests/golden/reference_kats.json, extracted from pkg/dealer/*_test.go) plus the derived vectors
of SURVEY.md §8c.  CPU only."""
import numpy as np
import pytest

import oracle as o
from tests import go_sort_py as gs

POL = {"binpack": o.POLICY_BINPACK, "spread": o.POLICY_SPREAD, "random": o.POLICY_RANDOM}


def test_binpack_choose_kats(kats):  # rater_test.go:133-266
    assert len(kats["binpack_choose"]["cases"]) == 5
    for c in kats["binpack_choose"]["cases"]:
        ok, idx, _ = o.choose(o.POLICY_BINPACK, c["free"], c["demand"])
        assert ok == (not c["expect_failed"])
        if ok:
            assert idx == c["expected"]


def test_spread_choose_kats(kats):  # rater_test.go:268-401
    assert len(kats["spread_choose"]["cases"]) == 5
    for c in kats["spread_choose"]["cases"]:
        ok, idx, _ = o.choose(o.POLICY_SPREAD, c["free"], c["demand"])
        assert ok == (not c["expect_failed"])
        if ok:
            assert idx == c["expected"]


def test_binpack_rate_kat(kats):  # rater_test.go:9-37
    for c in kats["binpack_rate"]["cases"]:
        assert (o.rate(o.POLICY_BINPACK, c["gpus1"]) < o.rate(o.POLICY_BINPACK, c["gpus2"])) == c["first_lower"]


def test_spread_rate_kats(kats):  # rater_test.go:39-131
    assert len(kats["spread_rate"]["cases"]) == 4
    for c in kats["spread_rate"]["cases"]:
        assert (o.rate(o.POLICY_SPREAD, c["gpus1"]) > o.rate(o.POLICY_SPREAD, c["gpus2"])) == c["first_preferred"]


def test_sample_choose_kats(kats):  # allocate_test.go:160-190
    assert len(kats["sample_choose"]["cases"]) == 4
    for c in kats["sample_choose"]["cases"]:
        ok, _, _ = o.choose(o.POLICY_RANDOM, c["free"], c["demand"])
        assert ok == c["success"]


def test_sortable_gpus_kat(kats):  # allocate_test.go:213-230
    for c in kats["sortable_gpus"]["cases"]:
        f, _, idx = o.go_sort(c["free"])
        assert f.tolist() == c["expected_free"] and idx.tolist() == c["expected_index"]


def test_gpu_resource_kats(kats):  # allocate_test.go:16-86 through GPUs.Allocate / Release on one cell
    g = kats["gpu_resource"]
    for c in g["sub"]:
        oc = o.OracleCluster(1)
        oc.node_set(0, [c["origin"]])
        assert oc.node_apply(0, [0], [c["target"]], +1) == 0
        assert oc.node_get(0)[1].tolist() == [c["expect"]]
    for c in g["add"]:
        oc = o.OracleCluster(1)
        oc.node_set(0, [c["origin"]])
        assert oc.node_apply(0, [0], [c["target"]], -1) == 0
        assert oc.node_get(0)[1].tolist() == [c["expect"]]
    for c in g["sub_if_available"]:  # CanAllocate gate: a demand that does not fit changes nothing
        oc = o.OracleCluster(1)
        oc.node_set(0, [c["origin"]])
        rc = oc.node_apply(0, [0], [c["target"]], +1)
        assert (rc == 0) == (c["origin"] >= c["target"])
        assert oc.node_get(0)[1].tolist() == [c["expect"]]


def test_derived_sort_rule_vectors(derived):
    for v in derived["sort_rule_choose"]:
        ok, idx, _ = o.choose(POL[v["policy"]], v["free"], v["demand"], rule=o.SORT_GO116)
        assert ok and idx == v["go116"]
        ok, idx, _ = o.choose(POL[v["policy"]], v["free"], v["demand"], rule=o.SORT_GO119)
        assert ok and idx == v["go119"]


def test_derived_rates(derived):
    for v in derived["rates"]:
        assert o.rate(POL[v["policy"]], v["free"]) == v["score"], v


def test_binpack_rate_is_fp64_not_integer():
    """int(float64(used)/float64(sum)*100) differs from used*100//sum exactly where fp64 rounds down"""
    diffs = []
    for used in range(0, 801):
        free = [100] * 8
        left = used
        for g in range(8):
            take = min(100, left)
            free[g] -= take
            left -= take
        r = o.rate(o.POLICY_BINPACK, free) + 8
        assert r == int(float(used) / float(800) * 100)
        if r != used * 100 // 800:
            diffs.append(used)
    assert diffs == [232, 456, 464]


@pytest.mark.parametrize("n", [2, 3, 6, 7, 8, 9, 12, 13, 20, 41, 64])
def test_go_sort_matches_python_restatement(n):
    rng = np.random.default_rng(1234 + n)
    for _ in range(300):
        free = rng.integers(0, 6, n) * 20  # many ties
        rl = rng.integers(0, 3, n)
        f, r, idx = o.go_sort(free, rl)
        cells = gs.Cells(free.tolist(), rl.tolist(), range(n))
        gs.go116_sort(cells)
        assert [c[2] for c in cells.a] == idx.tolist()
        keys = (f + 50 * r).tolist()
        assert keys == sorted(keys)


def test_choose_matches_python_restatement():
    rng = np.random.default_rng(7)
    for trial in range(3000):
        G = int(rng.integers(1, 13))
        C = int(rng.integers(1, 7))
        free = (rng.integers(0, 11, G) * 10).tolist()
        rl = rng.integers(0, 3, G).tolist() if trial % 2 else [0] * G
        demand = (rng.integers(0, 8, C) * 10).tolist()
        for spread in (False, True):
            ok, idx, ne = o.choose(o.POLICY_SPREAD if spread else o.POLICY_BINPACK, free, demand, rl=rl)
            pok, pidx = gs.choose(free, rl, demand, spread)
            assert ok == pok
            if ok:
                assert idx == pidx
            else:
                assert idx[:ne] == pidx


def test_go119_rule_differs_sometimes():
    rng = np.random.default_rng(3)
    differ = 0
    for _ in range(2000):
        free = (rng.integers(0, 6, 8) * 20).tolist()
        d = [int(rng.integers(1, 6) * 10)]
        a = o.choose(o.POLICY_SPREAD, free, d, rule=o.SORT_GO116)
        b = o.choose(o.POLICY_SPREAD, free, d, rule=o.SORT_GO119)
        assert a[0] == b[0]  # feasibility never depends on the tie order when RemainLoad is 0
        differ += a[1] != b[1]
    assert differ > 100


def _load_usage_py(values, ts, present, periods, now_ns, g):
    """GPUResource.LoadUsage (allocate.go:173-195) with GetUsage (nodeusage.go:82-111) and inUpdateTimePeriod
    (stats.go:30-55) folded in, restated independently of the C oracle.  values/ts/present: per metric."""
    import math

    usage = 0.0
    for metric, window_ns in periods:  # getActiveDuration already applied by the caller (stats.go:57-66)
        if window_ns == 0 or not present[metric]:  # activeDuration == 0 / node has no usage map: skipped
            continue
        if not (now_ns < int(ts[metric][g]) * 10**9 + window_ns):  # now.Before(updatetime)
            continue
        v = float(values[metric][g])
        if v < 0 or v > 1:
            continue
        usage += math.ceil(10 * v) / 10
    return usage, 2 - int(usage)


@pytest.mark.parametrize("periods", [
    [(0, 320 * 10**9), (1, 320 * 10**9)], [(1, 320 * 10**9), (0, 320 * 10**9)], [(0, 320 * 10**9)], [(1, 40 * 10**9), (1, 320 * 10**9), (0, 90 * 10**9)], []])
def test_load_aware_rates_match_python_restatement(periods):
    """the load-aware half of Binpack.Rate / Spread.Rate (rater.go:59-70,113-123) and the RemainLoad it leaves
    (allocate.go:193), oracle vs a second restatement in plain Python floats: same summation order, so sums that
    land just below an integer (0.1 * 10 = 0.9999999999999999) truncate the same way"""
    from nano_gpu_scheduler_b200 import workloads as wl

    now_s = 1_790_000_000
    now_ns = now_s * 10**9 + 987_654_321
    n = 400
    gcount, free = wl.make_cluster(n, 0.5, 0xFACE, hetero=True)
    tables = wl.make_load_tables(n, 0xFACE, now_s, stale_frac=0.2, absent_frac=0.1)
    oc = o.OracleCluster(n)
    oc.load_snapshot(gcount, free)
    for m, (value, ts, absent) in enumerate(tables):
        oc.load_update_bulk(m, value, ts, ok=np.repeat(~absent[:, None], 8, axis=1).astype(np.uint8))
    L = o.make_load_ctx(periods, now_ns)
    for policy in (o.POLICY_BINPACK, o.POLICY_SPREAD):
        got = oc.assume([0], policy, load=L)  # an empty-handed pod fits everywhere: the score is Rate alone
        _, _, rl = oc.snapshot()
        for i in range(n):
            G = int(gcount[i])
            values = [tables[0][0][i], tables[1][0][i]]
            tss = [tables[0][1][i], tables[1][1][i]]
            present = [not tables[0][2][i], not tables[1][2][i]]
            load, remain = 0.0, []
            for g in range(G):
                u, r = _load_usage_py(values, tss, present, periods, now_ns, g)
                load += u
                remain.append(r)
            fr = [int(v) for v in free[i][:G]]
            if policy == o.POLICY_BINPACK:
                want = int(float(100 * G - sum(fr)) / float(100 * G) * 100) + (int(load) // G) * 50 - G
            else:
                want = 100 * sum(1 for v in fr if v == 100) + sum(fr) // 10 - G - int(load) // G
            assert got["score"][i] == want, (policy, i)
            assert rl[i][:G].tolist() == remain, (policy, i)
    oc.close()


def test_go_vectors_when_present():
    """tests/golden/go_vectors.json — outputs of the reference's own Go dealer (tests/golden/go_vectors/gen_go_vectors.go,
    needs Go 1.16): tie order at 7..12 GPUs, Binpack.Rate over all totals, the load-aware mode.  Skipped while nobody with a
    Go toolchain has generated the file."""
    import json
    import os

    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "go_vectors.json")
    if not os.path.exists(path):
        pytest.skip("tests/golden/go_vectors.json has not been generated (needs Go 1.16 and the reference's module cache)")
    with open(path) as f:
        v = json.load(f)
    for c in v["choose"]:
        ok, idx, _ = o.choose(POL[c["policy"]], c["free"], c["demand"], rl=c["remain_load"])
        assert ok == c["ok"], c
        if ok:
            assert idx == c["plan"], c
    for c in v["binpack_rate"]:
        G, used = c["gpus"], c["used"]
        free, left = [], used
        for _ in range(G):
            take = min(left, 100)
            free.append(100 - take)
            left -= take
        assert o.rate(o.POLICY_BINPACK, free) == c["score"], c
    now_s = 1_790_000_000
    for c in v["load"]:
        G = len(c["free"])
        oc = o.OracleCluster(1)
        oc.node_set(0, c["free"])
        for metric, table in ((o.METRIC_CORE, c["core"]), (o.METRIC_MEMORY, c["memory"])):
            if table is None:
                continue  # the node has no entry: GetUsage says "not exist" (nodeusage.go:85-88)
            oc.load_add_node(0, metric)
            for card, s in enumerate(table):
                try:
                    val, val_ok = float(s["value"]), True
                except ValueError:
                    val, val_ok = 0.0, False
                oc.load_update(0, metric, card, val, now_s - s["age_s"], value_ok=val_ok, ts_ok=not s["bad_time"])
        # getActiveDuration (stats.go:57-66) takes the FIRST entry with the name; window = period + 5 min (type.go:6)
        first = {}
        for m, p in c["periods"]:
            first.setdefault(m, p)
        periods = [(m, first[m] + 5 * 60 * 10**9) for m, _ in c["periods"]]
        load = o.make_load_ctx(periods, now_s * 10**9 + 500_000_000)
        r = oc.assume(c["demand"], POL[c["policy"]], load=load, want_plan=True)
        assert bool(r["feasible"][0]) == c["ok"], c
        # (Plan.Score is Rate's value whether or not Choose succeeds; the oracle reports 0 for an infeasible node as
        # NodeInfo.Score does, node.go:59-68)
        if c["ok"]:
            assert r["score"][0] == c["score"], c
            assert r["plan"][0].tolist() == c["plan"], c
        _, _, rl = oc.node_get(0)
        assert rl.tolist()[:G] == c["remain_load_after"], c
