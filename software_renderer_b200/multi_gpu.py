"""Multi-GPU layer: independent views shard across ranks (one process per GPU); nothing is exchanged while
rendering.  A collective is used only when a composite is requested: finished framebuffers are gathered to rank 0
(NCCL over NVLink on the GPU box; the same code runs on gloo/CPU tensors for the host-logic tests).

A view depends only on (mesh, matrix, width, height) — Renderer.swift:204-230 keeps no cross-frame state — so the
partition is a plain block split of the view list.
"""
from __future__ import annotations

from typing import Tuple

import torch
import torch.distributed as dist


def shard_range(nviews: int, rank: int, world: int) -> Tuple[int, int]:
    """Views [lo, hi) rendered by `rank`: contiguous blocks, sizes differ by at most one, earlier ranks get the extra."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    base, rem = divmod(nviews, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def gather_framebuffers(local: torch.Tensor, nviews: int, dst: int = 0, group=None):
    """Gathers per-rank framebuffers (n_local, H, W, C) to `dst` in view order.  Returns the (nviews, H, W, C) tensor
    on `dst`, None elsewhere.  Shards may differ in size by one view, so each rank's block is padded to the largest
    shard for the collective and trimmed on arrival."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    sizes = [shard_range(nviews, r, world)[1] - shard_range(nviews, r, world)[0] for r in range(world)]
    assert local.shape[0] == sizes[rank], "local shard does not match shard_range"
    biggest = max(sizes)
    send = local
    if local.shape[0] < biggest:
        pad = torch.zeros((biggest - local.shape[0],) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
        send = torch.cat([local, pad], dim=0)
    send = send.contiguous()
    if rank == dst:
        recv = [torch.empty_like(send) for _ in range(world)]
        dist.gather(send, gather_list=recv, dst=dst, group=group)
        return torch.cat([recv[r][: sizes[r]] for r in range(world)], dim=0)
    dist.gather(send, gather_list=None, dst=dst, group=group)
    return None


def max_over_ranks(value: float, device=None, group=None) -> float:
    """Device-timed durations are reported as the max over ranks."""
    if not (dist.is_available() and dist.is_initialized()):
        return value
    t = torch.tensor([value], dtype=torch.float64, device=device if device is not None else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t.item())
