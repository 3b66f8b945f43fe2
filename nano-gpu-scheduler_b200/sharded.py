"""Node-range sharding over the GPUs of one box: one process and one GpuDealer per
GPU, launched under torchrun.  torch.distributed is plumbing only: it carries the 64-byte IPC
handles once at set-up and one small allreduce-max per batch that merges the owner's plan into
every rank's outputs.  The per-pod arg-max itself is traded between the ranks' kernels through
peer-mapped memory (ngs_dealer.cu, cross_rank_exchange)."""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist

from . import _abi
from .dealer import GpuDealer


def gather_handles(handle: bytes, group=None, device=None) -> bytes:
    """all-gather one fixed-size byte string per rank, rank-major"""
    world = dist.get_world_size(group)
    mine = torch.tensor(list(handle), dtype=torch.uint8, device=device)
    out = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(out, mine, group=group)
    return b"".join(bytes(t.cpu().tolist()) for t in out)


def merge_outputs(packed: torch.Tensor, group=None) -> torch.Tensor:
    """element-wise max over ranks of the packed [node | score | gpu_idx] int32 outputs: the owner
    of each pod holds the real values, every other rank holds values that are <= them
    (-1 / INT32_MIN fill, or the same node and score)."""
    dist.all_reduce(packed, op=dist.ReduceOp.MAX, group=group)
    return packed


class ShardedDealer:
    """rank r owns global nodes [r*nodes_per_rank, (r+1)*nodes_per_rank)"""

    def __init__(self, nodes_per_rank, rank=None, world=None, device=None, stream=None, group=None):
        self.rank = dist.get_rank(group) if rank is None else rank
        self.world = dist.get_world_size(group) if world is None else world
        self.group = group
        self.device = torch.cuda.current_device() if device is None else device
        self.n = int(nodes_per_rank)
        self.first = self.rank * self.n
        self.local = GpuDealer(self.n, device=self.device, first_node=self.first, stream=stream)
        if self.world > 1:
            handles = gather_handles(self.local.xchg_export(), group, torch.device("cuda", self.device))
            self.local.xchg_attach(self.rank, self.world, handles)
            dist.barrier(group)  # every rank has mapped every buffer before the first batch

    def close(self):
        self.local.close()

    def schedule_batch(self, percent, n_containers, policy, load=None, seed=0, seq0=0):
        """every rank passes the same pods; returns the merged (node, score, gpu_idx) on every rank"""
        pc = np.ascontiguousarray(percent, dtype=np.int32)
        P, max_c = pc.shape
        if P == 1:
            # one pod: the low-latency kernel trades keys (random policy: feasible counts) and the plan across the ranks
            # itself, no collective needed
            return self.local.schedule_batch(pc, n_containers, policy, load=load, seed=seed, seq0=seq0)
        dev = torch.device("cuda", self.device)
        t_pc = torch.from_numpy(pc).to(dev)
        t_nc = torch.from_numpy(np.ascontiguousarray(n_containers, dtype=np.int32)).to(dev)
        t_pol = torch.from_numpy(np.ascontiguousarray(np.broadcast_to(np.asarray(policy, dtype=np.int32), (P,)))).to(dev)
        out = self.schedule_batch_dev(P, t_pc, t_nc, max_c, t_pol, load=load, seed=seed, seq0=seq0)
        o = out.cpu().numpy()
        return o[:P].copy(), o[P:2 * P].copy(), o[2 * P:].reshape(P, max_c).copy()

    def schedule_batch_dev(self, P, t_pc, t_nc, max_c, t_pol, load=None, seed=0, seq0=0, out=None, same_stream=False):
        """device tensors in, packed int32 tensor [node(P) | score(P) | gpu_idx(P*max_c)] out"""
        if out is None:
            out = torch.empty((2 + max_c) * P, dtype=torch.int32, device=t_pc.device)
        base = out.data_ptr()
        self.local.schedule_batch_dev(P, t_pc.data_ptr(), t_nc.data_ptr(), max_c, t_pol.data_ptr(), base, base + 4 * P,
                                      base + 8 * P, load=load, seed=seed, seq0=seq0)
        if self.world > 1:
            if not same_stream:
                self.local.sync()  # the merge runs on torch's current stream; the batch ran on the handle's
            merge_outputs(out, self.group)
        return out


__all__ = ["ShardedDealer", "gather_handles", "merge_outputs", "_abi"]
