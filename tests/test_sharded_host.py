"""Host-side logic of the multi-GPU path on CPU: world_size-2 gloo groups exercise the handle
all-gather and the allreduce-max merge of per-rank outputs (what runs over NCCL on the box)."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

INT32_MIN = -(2**31)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from nano_gpu_scheduler_b200.sharded import gather_handles, merge_outputs

        # 1. handles: every rank ends up with all handles, rank-major
        mine = bytes([(rank * 37 + i) % 256 for i in range(64)])
        allh = gather_handles(mine)
        want = b"".join(bytes([(r * 37 + i) % 256 for i in range(64)]) for r in range(world))
        assert allh == want
        # 2. merge: pod p is owned by rank p % world; the others hold the fill values the kernel writes
        P, max_c, nodes_per_rank = 64, 3, 1000
        rng = np.random.default_rng(5)
        node = rng.integers(0, nodes_per_rank * world, P).astype(np.int32)
        node[::7] = -1  # unschedulable pods
        score = rng.integers(-8, 100, P).astype(np.int32)
        score[node < 0] = 0
        gidx = rng.integers(-1, 8, (P, max_c)).astype(np.int32)
        gidx[node < 0] = INT32_MIN
        owner = np.where(node >= 0, node // nodes_per_rank, 0)
        mine_g = np.where((owner == rank)[:, None], gidx, INT32_MIN).astype(np.int32)
        packed = torch.from_numpy(np.concatenate([node, score, mine_g.ravel()]).astype(np.int32))
        merged = merge_outputs(packed).numpy()
        assert np.array_equal(merged[:P], node) and np.array_equal(merged[P:2 * P], score)
        assert np.array_equal(merged[2 * P:].reshape(P, max_c), gidx)
        # random policy: non-owners report node -1 and the score of a feasible pick
        node_r = np.where(owner == rank, node, -1).astype(np.int32)
        packed = torch.from_numpy(np.concatenate([node_r, np.where(node >= 0, 100, 0).astype(np.int32), mine_g.ravel()]))
        merged = merge_outputs(packed).numpy()
        assert np.array_equal(merged[:P], node)
        q.put((rank, "ok"))
    except Exception as e:  # pragma: no cover
        q.put((rank, repr(e)))
    finally:
        dist.destroy_process_group()


def test_handle_gather_and_merge_world2():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    assert sorted(res) == [(0, "ok"), (1, "ok")], res
