"""The filter / prioritize routes end to end on the GPU: JSON body in -> C++ codec ->
name resolution -> ngs_assume / ngs_score on the device -> JSON body out, compared byte for byte with
what the reference would marshal for the ORACLE's verdicts on the same cluster (pkg/routes/routes.go:40-123,
pkg/scheduler/predicate.go:21-38, priority.go:30-37, failure texts of pkg/dealer/rater.go:47,99,152 and
dealer.go:102)."""
import json

import numpy as np
import pytest

import oracle as o
from nano_gpu_scheduler_b200.routes import Routes, node_name
from tests import extender_util as x

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", autouse=True)
def _built(built):
    import nano_gpu_scheduler_b200.build as b

    b.build_host()


def demand_str(d):
    return "[" + " ".join("(%d)" % v for v in d) + "]"


def expected_filter(oc, gcount, policy, demand, names):
    """what predicate.go builds from the oracle's Assume over `names` (unknown names -> the lister's error)"""
    ids = []
    for nm in names:
        ok = nm.startswith("node-") and len(nm) == 11 and nm[5:].isdigit() and int(nm[5:]) < len(gcount)
        ids.append(int(nm[5:]) if ok else -1)
    known = [i for i in ids if i >= 0]
    res = oc.assume(demand, policy, idx=np.array(known, dtype=np.uint32), want_plan=True) if known else None
    can, failed, k = [], {}, 0
    for nm, i in zip(names, ids):
        if i < 0:
            failed[nm] = 'nano gpu scheduler get node failed: node "%s" not found' % nm
            continue
        if res["feasible"][k]:
            can.append(nm)
        else:
            free = oc.node_get(i)[1][: gcount[i]]
            gpus = "".join("(%d)" % v for v in free)
            if policy == o.POLICY_RANDOM:
                failed[nm] = "unexpected failure of allocation  %s on %s" % (demand_str(demand), gpus)
            else:
                part = res["plan"][k][: res["n_emitted"][k]]
                failed[nm] = "can't allocate   %s on %s: indexes=[%s]" % (demand_str(demand), gpus, " ".join(str(int(v)) for v in part))
        k += 1
    return can, failed


def make_assumed(gcount, rng, frac):
    """random valid assumed pods; returns [(node, uid, percent, gpu_idx)]"""
    out = []
    for i in np.flatnonzero(rng.random(len(gcount)) < frac):
        free = [100] * int(gcount[i])
        for j in range(int(rng.integers(1, 4))):
            c = int(rng.integers(1, 4))
            pc, gi = [], []
            for _ in range(c):
                g = int(rng.integers(0, gcount[i]))
                p = int(5 * rng.integers(1, 13))
                if free[g] < p:
                    p = 0
                free[g] -= p
                pc.append(p)
                gi.append(g if p else -1)
            out.append((int(i), "pod-%d-%d" % (i, j), pc, gi))
    return out


@pytest.mark.parametrize("rater,policy", [(0, o.POLICY_BINPACK), (1, o.POLICY_SPREAD), (2, o.POLICY_RANDOM)])
def test_routes_match_the_oracle_byte_for_byte(rater, policy):
    rng = np.random.default_rng(100 + rater)
    n = 3000
    gcount = rng.integers(1, 9, size=n).astype(np.uint8)
    assumed = make_assumed(gcount, rng, 0.5)
    oc = o.OracleCluster(n)
    oc.load_snapshot(gcount, np.where(np.arange(8)[None, :] < gcount[:, None], 100, 0).astype(np.int32))
    for node, _, pc, gi in assumed:
        assert oc.node_apply(node, gi, pc, +1) == 0
    r = Routes(gcount, rater, assumed)
    try:
        names = [node_name(i) for i in range(n)]
        for demand in ([20, 40], [0, 100, 5], [60, 60, 60], [100] * 9):
            pod = x.pod_json(demand)
            body = json.dumps({"Pod": pod, "Nodes": None, "NodeNames": names})
            can, failed = expected_filter(oc, gcount, policy, demand, names)
            assert r.predicate(body) == x.go_filter_result(can, failed), demand
            out, panicked = r.prioritize(body)
            sc = oc.assume(demand, policy)["score"]
            assert not panicked and out == x.go_priorities(names, [int(v) for v in sc])
        c = r.counters()
        assert c["requests"] == 8 and c["cache_hits"] == 7  # one tokenise + resolve for eight requests over the same list
        # kube-scheduler's rotating start index, a subset, unknown names and a duplicate
        sub = names[1700:] + names[:400] + ["bogus", "node-999999", names[5], names[5], "node-<&>"]
        demand = [50, 50]
        body = json.dumps({"Pod": x.pod_json(demand), "NodeNames": sub})
        can, failed = expected_filter(oc, gcount, policy, demand, sub)
        got = r.predicate(body)
        assert got == x.go_filter_result(can, failed)
        back = json.loads(got)
        assert back["FailedNodes"]["bogus"] == 'nano gpu scheduler get node failed: node "bogus" not found'
        assert list(back["FailedNodes"]) == sorted(back["FailedNodes"], key=lambda s: s.encode())
        out, _ = r.prioritize(body)
        want = []
        for nm in sub:
            known = nm[5:].isdigit() and int(nm[5:]) < n
            want.append(int(oc.assume(demand, policy, idx=np.array([int(nm[5:])], dtype=np.uint32))["score"][0]) if known else 0)
        assert out == x.go_priorities(sub, want)
    finally:
        r.close()
        oc.close()


def test_routes_all_rows_fast_path_and_state_changes():
    n = 2048
    gcount = np.full(n, 8, dtype=np.uint8)
    oc = o.OracleCluster(n)
    oc.load_snapshot(gcount, np.full((n, 8), 100, dtype=np.int32))
    r = Routes(gcount, 0)
    try:
        names = [node_name(i) for i in range(n)]
        demand = [70, 40]
        body = json.dumps({"Pod": x.pod_json(demand), "NodeNames": names})
        # the first request materialises every node in list order -> the list IS rows 0..n-1
        assert r.predicate(body) == x.go_filter_result(names, {})
        assert r.counters()["all_rows"] == 1
        rng = np.random.default_rng(3)
        for k in range(80):  # fill some nodes through Dealer.Bind, mirrored on the oracle
            node = int(rng.integers(0, 4))
            rc, gi, err = r.bind(node_name(node), "b%d" % k, demand)
            orc, ogi, _ = oc.bind(node, demand, o.POLICY_BINPACK)
            assert (rc == 0) == (orc == 0), err
            if rc == 0:
                assert gi == [int(v) for v in ogi]
        can, failed = expected_filter(oc, gcount, o.POLICY_BINPACK, demand, names)
        assert len(failed) > 0
        assert r.predicate(body) == x.go_filter_result(can, failed)  # cached candidate set, fresh verdicts
        c = r.counters()
        assert c["cache_hits"] == 1 and c["all_rows"] == 2
    finally:
        r.close()
        oc.close()


def test_route_error_shapes():
    r = Routes(np.full(4, 8, dtype=np.uint8), 0)
    try:
        assert r.predicate(b"{") == x.go_filter_error("unexpected EOF")  # routes.go:54-60
        assert r.predicate(json.dumps({"Pod": x.pod_json([10]), "Nodes": {"items": []}})) == x.go_filter_error(
            "nano-gpu-scheduler extender must be configured with nodeCacheCapable=true")  # routes.go:62-68
        assert r.predicate(json.dumps({"Pod": x.pod_json([10]), "NodeNames": []})) == b'{"Nodes":null,"NodeNames":[],"FailedNodes":{},"Error":""}'
        out, panicked = r.prioritize(b"not json")
        assert panicked and out == b""  # routes.go:103: the reference panics here
        out, panicked = r.prioritize(json.dumps({"Pod": x.pod_json([10]), "NodeNames": []}))
        assert not panicked and out == b"[]"
    finally:
        r.close()


def test_status_route_body():
    """StatusRoute (routes.go:212-240) marshals map[string]*NodeInfo: node names in key order, each GPU's Percent /
    PercentTotal / RemainLoad; only nodes the dealer has built appear; PlanCache is written empty"""
    n = 40
    gcount = np.array([1 + i % 8 for i in range(n)], dtype=np.uint8)
    assumed = [(3, "a", [30, 20], [0, 2]), (17, "b", [100], [1])]
    oc = o.OracleCluster(n)
    oc.load_snapshot(gcount, np.where(np.arange(8)[None, :] < gcount[:, None], 100, 0).astype(np.int32))
    for node, _, pc, gi in assumed:
        assert oc.node_apply(node, gi, pc, +1) == 0
    with Routes(gcount, 1, assumed) as r:
        def want(nodes):
            items = []
            for i in sorted(nodes, key=node_name):
                _, free, rl = oc.node_get(i)
                gpus = ",".join('{"Percent":%d,"PercentTotal":100,"RemainLoad":%d}' % (free[g], rl[g]) for g in range(gcount[i]))
                items.append('"%s":{"Rater":{},"Name":"%s","GPUs":[%s],"PlanCache":{}}' % (node_name(i), node_name(i), gpus))
            return ("{" + ",".join(items) + "}").encode()

        assert r.status() == want([3, 17])  # NewDealer built the nodes that run assumed pods, nothing else
        names = [node_name(i) for i in (30, 5, 17, 8)]
        r.predicate(json.dumps({"Pod": x.pod_json([10]), "NodeNames": names}))
        assert r.status() == want([3, 17, 30, 5, 8])  # the filter call built the nodes it was asked about
        rc, gi, _ = r.bind(node_name(5), "c", [40, 40])
        orc, ogi, _ = oc.bind(5, [40, 40], o.POLICY_SPREAD)
        assert rc == 0 and orc == 0 and gi == [int(v) for v in ogi]
        got = r.status()
        assert got == want([3, 17, 30, 5, 8])
        assert json.loads(got)[node_name(5)]["GPUs"][0]["Percent"] in (60, 100)
    oc.close()


def test_bind_route_bodies():
    """BindRoute (routes.go:125-169) with NewNanoGPUBind (scheduler/bind.go:38-59): args in, {"Error": ...} out, 200 / 500"""
    n = 16
    gcount = np.full(n, 4, dtype=np.uint8)
    oc = o.OracleCluster(n)
    oc.load_snapshot(gcount, np.where(np.arange(8)[None, :] < 4, 100, 0).astype(np.int32).repeat(n, 0).reshape(n, 8))
    with Routes(gcount, 0) as r:
        r.add_pod("default", "p1", "uid-1", [60, 30])
        r.add_pod("default", "gone", "uid-2", [10], completed=True)
        r.add_pod("default", "big", "uid-3", [100, 100, 100, 100, 100])
        args = lambda name, uid, node: json.dumps({"PodName": name, "PodNamespace": "default", "PodUID": uid, "Node": node})  # noqa: E731
        body, st = r.bind_route(args("p1", "uid-1", node_name(2)))
        assert (body, st) == (b'{"Error":""}', 200)
        assert oc.bind(2, [60, 30], o.POLICY_BINPACK)[0] == 0
        got = json.loads(r.status())[node_name(2)]["GPUs"]
        assert [g["Percent"] for g in got] == [int(v) for v in oc.node_get(2)[1]]
        body, st = r.bind_route(args("nobody", "uid-9", node_name(2)))
        assert st == 500 and json.loads(body) == {"Error": 'pods "nobody" not found'}
        body, st = r.bind_route(args("p1", "other-uid", node_name(2)))
        assert st == 500 and json.loads(body) == {"Error": "pod p1 in ns default's uid is uid-1, and it's not equal with expected other-uid"}
        body, st = r.bind_route(args("gone", "uid-2", node_name(2)))
        assert st == 500 and json.loads(body) == {"Error": "pod gone/default already deleted or completed"}
        body, st = r.bind_route(args("big", "uid-3", node_name(3)))  # five full GPUs on a four-GPU node (node.go:75)
        assert st == 500 and json.loads(body) == {"Error": "assume [(100) (100) (100) (100) (100)] on (100)(100)(100)(100) failed"}
        body, st = r.bind_route(args("p1", "uid-1", "node-nowhere"))
        assert st == 500 and json.loads(body) == {"Error": 'node "node-nowhere" not found'}
        body, st = r.bind_route(b"{")
        assert st == 500 and body == b'{"Error":"unexpected EOF"}'
    oc.close()
