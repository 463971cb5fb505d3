"""Multi-GPU layer: independent views shard across ranks (one process per GPU); nothing is exchanged while
rendering.  A collective is used only when a composite is requested: finished framebuffers are gathered to rank 0
(NCCL over NVLink on the GPU box; the same code runs on gloo/CPU tensors for the host-logic tests).

A view depends only on (mesh, matrix, width, height) — Renderer.swift:204-230 keeps no cross-frame state — so the
partition is a plain block split of the view list.  One frame can also be split by rows (`band_range`, `gather_bands`
over `CudaRenderer.set_band` / `sr_set_band`): tiles are independent, each rank renders a band of tile rows.
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


def init_comm(renderer, group=None) -> None:
    """Joins `renderer`'s context to the job's NCCL communicator for sr_gather: rank 0 creates the NCCL id through the
    C ABI (sr_comm_unique_id), torch.distributed carries the 128 bytes to the other ranks (any backend: this is host
    plumbing), every rank calls sr_comm_init.  From here on the composite needs no torch on the data path."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    box = [type(renderer).comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0, group=group)
    renderer.comm_init(box[0], rank, world)


def view_block_bytes(nviews: int, world: int, bytes_per_view: int):
    """bytes_per_rank argument of sr_gather for a block split of `nviews` views."""
    return [(shard_range(nviews, r, world)[1] - shard_range(nviews, r, world)[0]) * bytes_per_view for r in range(world)]


def gather_views(renderer, local: torch.Tensor, nviews: int, out: torch.Tensor | None, slot: int = 0, dst: int = 0,
                 stream=None):
    """sr_gather of this rank's finished framebuffers (n_local, H, W[, C]) into `out` (nviews, ...) on `dst`, in view
    order, asynchronously on the context's gather stream (ordered after the rendering enqueued so far on `stream`).
    Call renderer.gather_sync() (host) or renderer.gather_fence(slot) (device) before touching the buffers again."""
    per_view = local[0].numel() * local.element_size() if local.shape[0] else (out[0].numel() * out.element_size())
    sizes = view_block_bytes(nviews, renderer.comm_world, per_view)
    assert local.is_contiguous() and (out is None or out.is_contiguous())
    renderer.gather(local.data_ptr() if local.numel() else 0, out.data_ptr() if out is not None else None, sizes,
                    root=dst, slot=slot, after_stream=stream)


def band_range(height: int, rank: int, world: int, align: int = 16) -> Tuple[int, int]:
    """Rows [y0, y1) of ONE frame rendered by `rank` (screen-band split): the frame's tile
    rows (16 pixel rows each, the kernels' tile height) are block-split like views, so every band but the last is
    a whole number of tiles.  A rank may get an empty band (y0 == y1) when there are fewer tile rows than ranks."""
    if height <= 0 or align <= 0:
        raise ValueError("bad height/align")
    t0, t1 = shard_range((height + align - 1) // align, rank, world)
    return min(t0 * align, height), min(t1 * align, height)


def gather_bands(local: torch.Tensor, height: int, dst: int = 0, group=None, align: int = 16):
    """Stitches the ranks' bands of one frame (or of n frames): `local` is (n, rows_r, W, C) or (rows_r, W, ...) with
    rows_r = this rank's band_range; returns the full (n, height, W, C) / (height, W, ...) image on `dst`, None
    elsewhere.  Bands differ in row count, so each is padded to the tallest for the collective and trimmed."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    spans = [band_range(height, r, world, align) for r in range(world)]
    rows = [b - a for a, b in spans]
    batched = local.dim() == 4
    x = local if batched else local.unsqueeze(0)
    assert x.shape[1] == rows[rank], "local band does not match band_range"
    tallest = max(rows)
    send = x
    if rows[rank] < tallest:
        pad = torch.zeros((x.shape[0], tallest - rows[rank]) + tuple(x.shape[2:]), dtype=x.dtype, device=x.device)
        send = torch.cat([x, pad], dim=1)
    send = send.contiguous()
    if rank == dst:
        recv = [torch.empty_like(send) for _ in range(world)]
        dist.gather(send, gather_list=recv, dst=dst, group=group)
        full = torch.cat([recv[r][:, : rows[r]] for r in range(world)], dim=1)
        return full if batched else full[0]
    dist.gather(send, gather_list=None, dst=dst, group=group)
    return None


def max_over_ranks(value: float, device=None, group=None) -> float:
    """Device-timed durations are reported as the max over ranks."""
    if not (dist.is_available() and dist.is_initialized()):
        return value
    t = torch.tensor([value], dtype=torch.float64, device=device if device is not None else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t.item())
