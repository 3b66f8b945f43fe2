"""Synthetic clusters and pod streams of BASELINE.json's configs (generators fixed in SURVEY.md
§8d).  One generator feeds the oracle, the GPU path and the bench, so both sides always see the
same snapshot.  PRNG: splitmix64 counter stream, seed 0xB2000000 + config number."""
from __future__ import annotations

import numpy as np

MASK = (1 << 64) - 1
EXTENDER_ACTIVE_NS = 5 * 60 * 10**9  # ExtenderAtivePeriod, /root/reference/pkg/dealer/type.go:6


def splitmix64(x):
    """vectorised splitmix64 finaliser over uint64 counters"""
    x = (np.asarray(x, dtype=np.uint64) + np.uint64(0x9E3779B97F4A7C15))
    x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return x ^ (x >> np.uint64(31))


class Stream:
    """counter-based stream: draw(n) returns n fresh uint64 values"""

    def __init__(self, seed):
        self.seed = np.uint64(seed & MASK)
        self.ctr = 0

    def draw(self, n):
        with np.errstate(over="ignore"):
            c = np.arange(self.ctr, self.ctr + n, dtype=np.uint64) + self.seed * np.uint64(0x2545F4914F6CDD1D)
            self.ctr += n
            return splitmix64(c)

    def randint(self, lo, hi, n):
        """uniform integers in [lo, hi] (modulo bias is irrelevant for synthetic data)"""
        return (self.draw(n) % np.uint64(hi - lo + 1)).astype(np.int64) + lo

    def uniform(self, n):
        return (self.draw(n) >> np.uint64(11)).astype(np.float64) / float(1 << 53)


def make_cluster(n_nodes, p_used, seed, gpus_per_node=8, hetero=False):
    """free[n][8]: with probability p_used a cell's free = 100 - 5*U{1..20}, else 100."""
    st = Stream(seed)
    used = st.uniform(n_nodes * 8) < p_used
    amount = 5 * st.randint(1, 20, n_nodes * 8)
    free = np.where(used, 100 - amount, 100).astype(np.int32).reshape(n_nodes, 8)
    if hetero:
        gcount = st.randint(1, 8, n_nodes).astype(np.uint8)
    else:
        gcount = np.full(n_nodes, gpus_per_node, dtype=np.uint8)
    free[np.arange(8)[None, :] >= gcount[:, None]] = 0
    return gcount, free


def make_pods(n_pods, c_lo, c_hi, pct_lo, pct_hi, seed, max_c=None):
    """percent[p][max_c] with C = U{c_lo..c_hi} containers of 5*U{pct_lo..pct_hi} percent."""
    st = Stream(seed ^ 0x5EED)
    max_c = max_c or c_hi
    nc = st.randint(c_lo, c_hi, n_pods).astype(np.int32)
    pc = (5 * st.randint(pct_lo, pct_hi, n_pods * max_c)).astype(np.int32).reshape(n_pods, max_c)
    pc[np.arange(max_c)[None, :] >= nc[:, None]] = 0
    return pc, nc


def make_load_tables(n_nodes, seed, now_s, stale_frac=0.05, absent_frac=0.02):
    """per metric: value[n][8] in {0.0 .. 1.0 step 0.1} (+ a little noise so ceil matters),
    ts[n][8] (stale_frac of them older than any window), absent[n] nodes with no table entry."""
    st = Stream(seed ^ 0x10AD)
    out = []
    for _m in range(2):
        tenth = st.randint(0, 10, n_nodes * 8).astype(np.float64)
        jitter = st.randint(0, 9999, n_nodes * 8).astype(np.float64) / 1e5  # %.5f granularity
        value = np.clip(np.round(tenth / 10.0 - jitter * (tenth > 0), 5), 0.0, 1.0).reshape(n_nodes, 8)
        age = st.randint(0, 200, n_nodes * 8)
        stale = st.uniform(n_nodes * 8) < stale_frac
        ts = (now_s - np.where(stale, 100000 + age, age)).astype(np.int64).reshape(n_nodes, 8)
        absent = st.uniform(n_nodes) < absent_frac
        out.append((value, ts, absent))
    return out


DEFAULT_PERIODS = [(0, 15 * 10**9 + EXTENDER_ACTIVE_NS), (1, 15 * 10**9 + EXTENDER_ACTIVE_NS)]
# deploy/dynamic-scheduler-node-annotator-cm.yaml:11-16: core 15s, memory 15s

CONFIGS = {
    # name: nodes, p_used, pods, (c_lo, c_hi), (pct_lo, pct_hi), policy, load
    "C1": dict(nodes=10, p_used=0.0, pods=100, c=(1, 1), pct=(1, 20), policy=0, load=False, seed=0xB2000001),
    "C2": dict(nodes=5000, p_used=0.3, pods=10000, c=(1, 1), pct=(1, 20), policy=1, load=False, seed=0xB2000002),
    "C3": dict(nodes=100000, p_used=0.5, pods=None, c=(2, 4), pct=(0, 12), policy=0, load=True, seed=0xB2000003),
    "C4": dict(nodes=1000000, p_used=0.5, pods=None, c=(1, 1), pct=(1, 20), policy=0, load=False, seed=0xB2000004),
    "C5": dict(nodes=50000, p_used=0.5, pods=100000, c=(1, 3), pct=(1, 12), policy=None, load=False, seed=0xB2000005),
}
