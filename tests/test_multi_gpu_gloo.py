"""Host-side logic of the N>1 path on CPU: world_size-2 gloo processes shard a view list, "render" their shard with the
oracle (the CUDA library needs a GPU), and gather the framebuffers to rank 0 exactly as bench.py / the GPU box does
with NCCL.  The gathered composite must equal the single-process result."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from software_renderer_b200 import scene as S  # noqa: E402
from software_renderer_b200.multi_gpu import band_range, gather_bands, gather_framebuffers, max_over_ranks, shard_range  # noqa: E402

NVIEWS, W, H = 5, 96, 64


def test_shard_range_partitions_exactly():
    for n in (0, 1, 5, 8, 4096, 4097):
        for world in (1, 2, 3, 4, 8):
            spans = [shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            for a, b in zip(spans, spans[1:]):
                assert a[1] == b[0]
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_range(4, 2, 2)


def _render_views(views):
    from oracle import oracle as O
    v, i = S.icosphere(2)
    out = np.empty((len(views), H, W, 4), np.uint8)
    for k, m in enumerate(views):
        rc, c, _, _ = O.render(v, i, m, W, H)
        assert rc == 0
        out[k] = c
    return out


def _worker(rank, world, port, result_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        views = S.fibonacci_views(NVIEWS)
        lo, hi = shard_range(NVIEWS, rank, world)
        local = torch.from_numpy(_render_views(views[lo:hi]))
        full = gather_framebuffers(local, NVIEWS, dst=0)
        t = max_over_ranks(float(rank + 1))
        assert t == float(world)
        if rank == 0:
            np.save(result_path, full.numpy())
        else:
            assert full is None
    finally:
        dist.destroy_process_group()


def test_sharded_views_gather_equals_single_process(tmp_path):
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    path = str(tmp_path / "composite.npy")
    mp.spawn(_worker, args=(2, port, path), nprocs=2, join=True)
    got = np.load(path)
    want = _render_views(S.fibonacci_views(NVIEWS))
    assert got.shape == want.shape and np.array_equal(got, want)


def test_band_range_partitions_rows_on_tile_boundaries():
    for height in (1, 15, 16, 17, 200, 1080, 2160):
        for world in (1, 2, 3, 4, 8):
            spans = [band_range(height, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == height
            for a, b in zip(spans, spans[1:]):
                assert a[1] == b[0] and a[0] <= a[1]
            assert all(y0 % 16 == 0 for y0, _ in spans if y0 < height)
            tiles = [(y1 - y0 + 15) // 16 for y0, y1 in spans]
            assert max(tiles) - min(tiles) <= 1
    assert band_range(1080, 7, 8) == (960, 1080)       # 68 tile rows over 8 ranks: 9,9,9,9,8,8,8,8; the last is partial
    with pytest.raises(ValueError):
        band_range(0, 0, 1)


def _band_worker(rank, world, port, result_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        frames = _render_views(S.fibonacci_views(2))                 # (2, H, W, 4): every rank could render the frame;
        y0, y1 = band_range(H, rank, world)                          # it keeps only its band, as sr_set_band would
        full = gather_bands(torch.from_numpy(frames[:, y0:y1].copy()), H, dst=0)
        one = gather_bands(torch.from_numpy(frames[0, y0:y1].copy()), H, dst=0)
        if rank == 0:
            np.save(result_path, np.concatenate([full.numpy(), one.numpy()[None]], axis=0))
        else:
            assert full is None and one is None
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_banded_frame_gather_equals_whole_frame(tmp_path, world):
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    path = str(tmp_path / "bands.npy")
    mp.spawn(_band_worker, args=(world, port, path), nprocs=world, join=True)
    got = np.load(path)
    want = _render_views(S.fibonacci_views(2))
    assert np.array_equal(got[:2], want) and np.array_equal(got[2], want[0])


# ---- round 2: the host plumbing of sr_gather (the NCCL calls themselves need GPUs: tests/test_gather_gpu.py) ----------
from software_renderer_b200.multi_gpu import init_comm, view_block_bytes  # noqa: E402


def test_view_block_bytes_and_chunked_layout_is_view_order():
    """bench.py / nccl_composite_check.py gather chunk by chunk: chunk k = views [kC, (k+1)C), split over the ranks by
    shard_range; sr_gather lays rank r's block at the sum of the blocks before it, which must be view order."""
    for nviews, world, per in ((4096, 8, 128), (37, 2, 5), (10, 4, 3), (5, 8, 1)):
        C = per * world
        order = []
        for k in range((nviews + C - 1) // C):
            base, n_all = k * C, min(C, nviews - k * C)
            sizes = view_block_bytes(n_all, world, 7)
            assert sum(sizes) == 7 * n_all and len(sizes) == world
            for r in range(world):
                lo, hi = shard_range(n_all, r, world)
                assert sizes[r] == 7 * (hi - lo) and hi - lo <= per
                order += list(range(base + lo, base + hi))
        assert order == list(range(nviews))


class _FakeRenderer:
    """Stands in for CudaRenderer in init_comm: the id comes from 'rank 0's library', every rank must receive it."""
    made = 0

    @staticmethod
    def comm_unique_id():
        _FakeRenderer.made += 1
        return bytes(range(128))

    def comm_init(self, uid, rank, world):
        self.args = (uid, rank, world)


def _comm_worker(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        r = _FakeRenderer()
        init_comm(r)
        assert r.args == (bytes(range(128)), rank, world)
        assert _FakeRenderer.made == (1 if rank == 0 else 0)       # only the root asks the library for an id
        open(os.path.join(out_dir, f"ok{rank}"), "w").close()
    finally:
        dist.destroy_process_group()


def test_init_comm_hands_rank0s_id_to_every_rank(tmp_path):
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    mp.spawn(_comm_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert os.path.exists(tmp_path / "ok0") and os.path.exists(tmp_path / "ok1")
