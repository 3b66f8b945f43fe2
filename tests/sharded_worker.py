"""Run under torchrun (one rank per GPU): the node-range sharded path against the CPU oracle on
the whole cluster.  Rank 0 prints SHARDED OK on success."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import nano_gpu_scheduler_b200 as ngs  # noqa: E402
import oracle as o  # noqa: E402
from nano_gpu_scheduler_b200 import workloads as wl  # noqa: E402
from nano_gpu_scheduler_b200.sharded import ShardedDealer  # noqa: E402
from tests.parity_util import NOW_NS, NOW_S  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 3000  # nodes per rank
    N = n * world
    gcount, free = wl.make_cluster(N, 0.5, 0x5A4D)
    tables = wl.make_load_tables(N, 0x5A4D, NOW_S)
    sd = ShardedDealer(n, device=local)
    lo, hi = rank * n, (rank + 1) * n
    sd.local.snapshot_load(gcount[lo:hi], free[lo:hi], first=lo)
    for m, (value, ts, absent) in enumerate(tables):
        ok = np.repeat(~absent[lo:hi, None], 8, axis=1).astype(np.uint8)
        sd.local.load_update_bulk(m, value[lo:hi], ts[lo:hi], value_ok=ok, ts_ok=ok, first=lo)
    oc = None
    if rank == 0:
        oc = o.OracleCluster(N)
        oc.load_snapshot(gcount, free)
        for m, (value, ts, absent) in enumerate(tables):
            oc.load_update_bulk(m, value, ts, ok=np.repeat(~absent[:, None], 8, axis=1).astype(np.uint8))
    # "wide": every 17th node has 9..12 GPUs (cells 8..11 in the second table plane, one-thread evaluation)
    wide = len(sys.argv) > 2 and sys.argv[2] == "wide"
    wide_rows = []
    if wide:
        wrng = np.random.default_rng(0x12)
        for i in range(3, N, 17):
            G = int(wrng.integers(9, 13))
            fr = np.where(wrng.random(G) < 0.5, wrng.integers(0, 21, G) * 5, 100).astype(np.int32)
            wide_rows.append(i)
            if lo <= i < hi:
                sd.local.node_set_wide(i, G, fr)
            if rank == 0:
                oc.node_set(i, fr)
    failures = 0
    seq = 0
    for rnd, (policy, load_on, P) in enumerate([(0, False, 300), (1, False, 300), (2, False, 200), (0, True, 300), (None, False, 300),
                                                (1, True, 100)]):
        pc, nc = wl.make_pods(P, 1, 4, 0, 12, 0x77 + rnd, max_c=4)
        pol = np.full(P, policy, dtype=np.int32) if policy is not None else (np.arange(P) % 3).astype(np.int32)
        gl = ngs.make_load_ctx(wl.DEFAULT_PERIODS, NOW_NS) if load_on else None
        node, score, gidx = sd.schedule_batch(pc, nc, pol, load=gl, seed=42, seq0=seq)
        if rank == 0:
            ol = o.make_load_ctx(wl.DEFAULT_PERIODS, NOW_NS) if load_on else None
            onode, oscore, ogidx = oc.schedule_batch(pc, nc, pol, load=ol, seed=42, seq0=seq, n_threads=4)
            good = np.array_equal(node, onode) and np.array_equal(score, oscore) and np.array_equal(gidx, ogidx)
            if not good:
                failures += 1
                bad = np.flatnonzero((node != onode) | (score != oscore))
                print(f"round {rnd} policy {policy} load {load_on}: MISMATCH at pods {bad[:8]}", flush=True)
        seq += P
    # single pods: the low-latency path (K7 with the cross-GPU trade and the plan mailbox), every rank gets the whole decision
    for rnd, (policy, load_on) in enumerate([(0, False), (1, True), (0, True), (2, False)]):
        pc, nc = wl.make_pods(60, 1, 4, 0, 12, 0x99 + rnd, max_c=4)
        gl = ngs.make_load_ctx(wl.DEFAULT_PERIODS, NOW_NS) if load_on else None
        ol = o.make_load_ctx(wl.DEFAULT_PERIODS, NOW_NS) if load_on else None
        launches0 = sd.local.stats()["kernel_launches"]
        for p in range(60):
            node, score, gidx = sd.schedule_batch(pc[p:p + 1], nc[p:p + 1], policy, load=gl, seed=42, seq0=seq)
            want = None
            if rank == 0:
                want = oc.schedule_batch(pc[p:p + 1], nc[p:p + 1], policy, load=ol, seed=42, seq0=seq)
                if not (np.array_equal(node, want[0]) and np.array_equal(score, want[1]) and np.array_equal(gidx, want[2])):
                    failures += 1
                    print(f"single pod {p} policy {policy} load {load_on}: MISMATCH {node} {score} {gidx} vs {want}", flush=True)
            # every rank must hold the same full answer: compare with rank 0's
            mine = torch.from_numpy(np.concatenate([node, score, gidx.ravel()]).astype(np.int32)).cuda()
            ref = mine.clone()
            dist.broadcast(ref, 0)
            if not torch.equal(mine, ref):
                failures += 1
                print(f"rank {rank}: single pod {p} differs from rank 0", flush=True)
            seq += 1
        if sd.local.stats()["kernel_launches"] - launches0 != 60:
            failures += 1
            print(f"rank {rank}: the single-pod kernel did not run (policy {policy})", flush=True)
    fl = torch.tensor([failures], device="cuda")
    dist.all_reduce(fl)
    failures = int(fl.item())
    # the shards together hold exactly the oracle's snapshot
    _, fr, rl = sd.local.snapshot_read(first=lo, n=n)
    t = torch.from_numpy(fr.astype(np.int32)).cuda()
    allfr = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(allfr, t)
    if rank == 0:
        got = torch.cat(allfr).cpu().numpy()
        _, ofr, _ = oc.snapshot()
        if not np.array_equal(got, ofr):
            failures += 1
            print("snapshot MISMATCH", flush=True)
    if wide:  # ... all twelve cells of the wide nodes
        mine = np.zeros((len(wide_rows), 12), dtype=np.int32)
        for k, i in enumerate(wide_rows):
            if lo <= i < hi:
                mine[k] = sd.local.node_get_wide(i)[1]
        t = torch.from_numpy(mine).cuda()
        dist.all_reduce(t)  # every row is non-zero on its owner only
        if rank == 0:
            got = t.cpu().numpy()
            for k, i in enumerate(wide_rows):
                G, ofr, _ = oc.node_get(i)
                if got[k, :G].tolist() != ofr.tolist():
                    failures += 1
                    print(f"wide node {i}: {got[k]} vs {ofr}", flush=True)
                    break
    if rank == 0:
        print("SHARDED OK" if failures == 0 else f"SHARDED FAILED ({failures})", flush=True)
    sd.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
