"""Where the wall time of one filter / prioritize request goes at 100k candidate names:
JSON decode, name resolution, the dealer call (H2D + scan + D2H + failure read-back) and JSON encode.
Run on the GPU box:  python profiles/tools/route_bench.py [n_nodes] > gpurun_out/route_bench.json"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from tests import extender_util as x  # noqa: E402
from nano_gpu_scheduler_b200.routes import Routes, node_name  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000
    gcount = np.full(n, 8, dtype=np.uint8)
    # every second node has all eight GPUs at 40 % free: a [50] demand fails there, a [10] demand fits everywhere
    assumed = [(i, "p%d" % i, [60] * 8, list(range(8))) for i in range(0, n, 2)]
    r = Routes(gcount, 0, assumed)
    names = [node_name(i) for i in range(n)]
    rot = names[n // 3:] + names[: n // 3]
    out = {"n_nodes": n, "restart_rebuild_ms": round(r.start_ms, 2), "restart_rebuild_nodes": len(assumed), "cases": []}

    def run(label, route, body, reps):
        body = body.encode()
        r.predicate(body) if route == "filter" else r.prioritize(body)  # warm (and fill the cache for the cached cases)
        c0 = r.counters()
        t = []
        for _ in range(reps):
            a = time.perf_counter()
            res = r.predicate(body) if route == "filter" else r.prioritize(body)[0]
            t.append(time.perf_counter() - a)
        c1 = r.counters()
        k = float(reps) * 1e6
        out["cases"].append({
            "case": label, "route": route, "body_mb": round(len(body) / 1e6, 2), "answer_mb": round(len(res) / 1e6, 2),
            "ms_p50": round(float(np.median(t)) * 1e3, 3), "ms_min": round(min(t) * 1e3, 3),
            "ms_decode": round((c1["ns_decode"] - c0["ns_decode"]) / k, 3), "ms_resolve": round((c1["ns_resolve"] - c0["ns_resolve"]) / k, 3),
            "ms_dealer": round((c1["ns_dealer"] - c0["ns_dealer"]) / k, 3), "ms_encode": round((c1["ns_encode"] - c0["ns_encode"]) / k, 3),
            "cache_hits": c1["cache_hits"] - c0["cache_hits"], "all_rows": c1["all_rows"] - c0["all_rows"]})

    fits = json.dumps({"Pod": x.pod_json([10]), "NodeNames": names})
    half = json.dumps({"Pod": x.pod_json([50]), "NodeNames": names})
    run("cached list, every node fits", "filter", fits, 20)
    run("cached list, half the nodes fail (50k messages)", "filter", half, 10)
    run("cached list", "prioritize", fits, 20)
    # alternate two different lists so that every request misses the cache: tokenise + 100k hash lookups each time
    a = json.dumps({"Pod": x.pod_json([10]), "NodeNames": rot})
    t = []
    c0 = r.counters()
    for i in range(10):
        body = (a if i % 2 == 0 else fits).encode()
        s = time.perf_counter()
        r.predicate(body)
        t.append(time.perf_counter() - s)
    c1 = r.counters()
    out["cases"].append({"case": "cache miss every request (rotated list), every node fits", "route": "filter",
                         "ms_p50": round(float(np.median(t)) * 1e3, 3), "ms_decode": round((c1["ns_decode"] - c0["ns_decode"]) / 1e7, 3),
                         "ms_resolve": round((c1["ns_resolve"] - c0["ns_resolve"]) / 1e7, 3), "ms_dealer": round((c1["ns_dealer"] - c0["ns_dealer"]) / 1e7, 3),
                         "ms_encode": round((c1["ns_encode"] - c0["ns_encode"]) / 1e7, 3), "cache_hits": c1["cache_hits"] - c0["cache_hits"]})
    # Python's own json for scale (C implementation): decode + encode of the same bodies
    s = time.perf_counter()
    obj = json.loads(fits)
    out["python_json_loads_ms"] = round((time.perf_counter() - s) * 1e3, 2)
    s = time.perf_counter()
    json.dumps({"Nodes": None, "NodeNames": obj["NodeNames"], "FailedNodes": {}, "Error": ""})
    out["python_json_dumps_ms"] = round((time.perf_counter() - s) * 1e3, 2)
    r.close()
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
