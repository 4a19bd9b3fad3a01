"""Node sharding across GPUs: one process per GPU, contiguous row ranges, one exchange.

Nodes are independent and a pod's answer is an argmax over nodes, so every rank
scores its own row range with no data-path collective; the only exchange is an
all-gather of the per-pod candidate keys (8 bytes per pod per rank) followed by the
merge kernel (K2).  The key = (order-preserving score bits << 32) | global row makes
"max over shards" identical to the single-GPU canonical winner (ties -> highest row).
torch.distributed is plumbing only (NCCL on GPUs; gloo for the CPU tests of this
host logic).
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import numpy as np


def shard_range(n_nodes: int, world: int, rank: int) -> Tuple[int, int]:
    """rows [lo, hi) of `rank`: ceil(N/G) rows each, the last shards may be short/empty"""
    per = (n_nodes + world - 1) // world
    lo = min(rank * per, n_nodes)
    return lo, min(lo + per, n_nodes)


def order_bits(score: np.ndarray) -> np.ndarray:
    """monotone float32 -> uint32 map used inside the candidate key (vgs_layout.h)"""
    b = np.asarray(score, dtype=np.float32).view(np.uint32)
    neg = (b >> np.uint32(31)).astype(bool)
    return np.where(neg, ~b, b | np.uint32(0x80000000)).astype(np.uint32)


def unorder_bits(ob: np.ndarray) -> np.ndarray:
    ob = np.asarray(ob, dtype=np.uint32)
    pos = (ob >> np.uint32(31)).astype(bool)
    return np.where(pos, ob ^ np.uint32(0x80000000), ~ob).astype(np.uint32).view(np.float32)


def encode_keys(best_row: np.ndarray, score: np.ndarray) -> np.ndarray:
    """candidate keys of a shard result (global rows); 0 where best_row < 0"""
    low = best_row.astype(np.int64).astype(np.uint64) & np.uint64(0xFFFFFFFF)
    k = (order_bits(score).astype(np.uint64) << np.uint64(32)) | low
    return np.where(best_row >= 0, k, np.uint64(0)).astype(np.uint64)


def decode_keys(keys: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
    keys = np.asarray(keys, dtype=np.uint64)
    row = np.where(keys != 0, (keys & np.uint64(0xFFFFFFFF)).astype(np.int64), -1).astype(np.int32)
    score = np.where(keys != 0, unorder_bits((keys >> np.uint64(32)).astype(np.uint32)), np.float32(0)).astype(np.float32)
    return row, score


def allgather_merge(keys: np.ndarray, counts: np.ndarray, group=None) -> Tuple[np.ndarray, np.ndarray]:
    """host-tensor version of the exchange (gloo): all-gather keys/counts, max / sum.
    The uint64 keys travel as int64 bit patterns; max is taken on the unsigned view."""
    import torch
    import torch.distributed as dist

    world = dist.get_world_size(group)
    k = torch.from_numpy(np.ascontiguousarray(keys).view(np.int64).copy())
    c = torch.from_numpy(np.ascontiguousarray(counts, dtype=np.int32).copy())
    gk = [torch.empty_like(k) for _ in range(world)]
    gc = [torch.empty_like(c) for _ in range(world)]
    dist.all_gather(gk, k, group=group)
    dist.all_gather(gc, c, group=group)
    allk = np.stack([t.numpy().view(np.uint64) for t in gk])
    allc = np.stack([t.numpy() for t in gc])
    return allk.max(axis=0), allc.sum(axis=0).astype(np.int32)


class _DevMem:
    """expose a raw device pointer to torch through __cuda_array_interface__"""

    def __init__(self, ptr: int, n: int, typestr: str):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False), "version": 3,
                                         "strides": None}


def device_view(ptr: int, n: int, dtype, device):
    import torch

    typestr = {torch.int64: "<i8", torch.int32: "<i4"}[dtype]
    return torch.as_tensor(_DevMem(ptr, n, typestr), device=device)


class ShardedScorer:
    """One rank of the node-sharded scoring pass (rank = GPU).  Mirrors what the Go
    extender would do with one context per GPU inside a single process."""

    def __init__(self, n_nodes_total: int, dmax: int, max_pods: int, max_reqs: Optional[int] = None,
                 max_picks: Optional[int] = None, device: int = 0, rank: int = 0, world: int = 1, group=None,
                 tuning=None):
        import torch

        from .capi import Context

        self.rank, self.world, self.group = rank, world, group
        max_pods += max_pods & 1  # the 12-byte-per-pod candidate block is exchanged in 8-byte units
        self.lo, self.hi = shard_range(n_nodes_total, world, rank)
        self.device = torch.device("cuda", device)
        torch.cuda.set_device(self.device)
        # a stream of our own, shared by the library, the collectives and the caller's timing
        # events: run calls under `with torch.cuda.stream(scorer.stream)`
        self.stream = torch.cuda.Stream(self.device)
        stream = self.stream.cuda_stream
        assert stream != 0
        self.ctx = Context(max_nodes=max(self.hi - self.lo, 1), max_devices=dmax, max_pods=max_pods,
                           max_reqs=max_reqs, max_picks=max_picks, device=device, row_base=self.lo, stream=stream,
                           tuning=tuning)
        self.max_pods = max_pods
        self.max_picks = self.ctx.cfg.max_picks
        # exchange block = [keys | counts | picks] in one allocation: ONE all-gather per pass moves every
        # shard's candidates together with the picks of its own best nodes
        bp, nbytes = self.ctx.exchange_block()
        assert nbytes % 4 == 0
        self.block_bytes = nbytes
        self.block = device_view(bp, nbytes // 4, torch.int32, self.device)
        if world > 1:
            self.gathered = torch.empty((world, nbytes // 4), dtype=torch.int32, device=self.device)

    def upload(self, snap_shard):
        assert snap_shard.n_nodes == self.hi - self.lo
        self.ctx.upload(snap_shard)

    def step_resident(self, n_picks: int):
        """score -> exchange -> merge -> resolve, everything on the device (self.stream)"""
        import torch
        import torch.distributed as dist

        with torch.cuda.stream(self.stream):
            self.ctx.score_resident()
            if self.world == 1:
                self.ctx.resolve_resident()
                return
            # every shard resolves its own best node per pod, the blocks travel once, and the merge keeps
            # the winning shard's result and picks (vgs_exchange_merge)
            self.ctx.resolve_local()
            dist.all_gather_into_tensor(self.gathered, self.block, group=self.group)
            self.ctx.exchange_merge(self.gathered.data_ptr(), self.world, self.block_bytes)

    def score(self, pods: np.ndarray, reqs: np.ndarray):
        self.ctx.pods_upload(pods, reqs)
        self.step_resident(int(reqs["nums"].sum()) if len(reqs) else 0)
        return self.ctx.results_download()

    def close(self):
        self.ctx.close()
