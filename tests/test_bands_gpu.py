"""Screen-band split of one frame (sr_set_band): rendering rows [y0, y1) on their own must give
exactly those rows of the whole frame — bit for bit against the CPU oracle's whole-frame image — for any cut
(tile-aligned or not), both visibility modes, every primitive type and every host policy.  The reference has no such
call: a band is defined by the frame Renderer.render(renderPass:) (Renderer.swift:204-230) produces."""
import os

import numpy as np
import pytest

import software_renderer_b200 as srb
from oracle import oracle as O
from software_renderer_b200 import scene as S
from software_renderer_b200.multi_gpu import band_range

from test_fuzz_gpu import POLICIES, scene as fuzz_scene

pytestmark = pytest.mark.gpu


def cuts(rng, H):
    """A random partition of [0, H) into 1..5 bands: arbitrary rows, plus tile-aligned and one-row bands."""
    n = int(rng.integers(0, 5))
    ys = sorted(set(int(y) for y in rng.integers(1, max(H, 2), size=n) if 0 < y < H))
    if H > 40 and rng.random() < 0.5:
        ys = sorted(set(ys + [16 * int(rng.integers(1, H // 16))]))
    if H > 3 and rng.random() < 0.3:
        y = int(rng.integers(1, H - 1))
        ys = sorted(set(ys + [y, y + 1]))
    edges = [0] + ys + [H]
    return list(zip(edges[:-1], edges[1:]))


def render_in_bands(r, v, i, m, W, H, bands, prim=srb.PrimitiveType.triangle, depth_test=False):
    """Whole-size images filled with a stale pattern; every band call may only write its own rows."""
    r.depth_test = depth_test
    rp = srb.RenderPass(srb.Image.color(W, H), srb.Image.depth(W, H), v, i, prim, m)
    rp.color_buffer.array[:] = 0xC3
    rp.depth_buffer.array[:] = 5.0
    try:
        for y0, y1 in bands:
            before_c = rp.color_buffer.array.copy()
            r.set_band(y0, y1)
            try:
                r.render(rp)
            except srb.DomainError:
                pass
            untouched = np.ones(H, bool)
            untouched[y0:y1] = False
            assert np.array_equal(rp.color_buffer.array[untouched], before_c[untouched]), f"band {y0}:{y1} wrote outside its rows"
    finally:
        r.set_band()
        r.depth_test = False
    return rp.color_buffer.array, rp.depth_buffer.array


@pytest.fixture(scope="module")
def cases():
    rng = np.random.default_rng(0xBA2D)
    out = []
    for k in range(20):
        v, i, m, W, H = fuzz_scene(rng, dense=k in (5, 13))
        want = {}
        for dt in (False, True):
            rc, rc_c, rc_d, _ = O.render(v, i, m, W, H, flags=O.DEPTH_LESS if dt else 0)
            assert rc == O.OK
            want[dt] = (rc_c, rc_d)
        out.append((v, i, m, W, H, cuts(rng, H), want))
    return out


@pytest.mark.parametrize("policy", POLICIES, ids=lambda p: ",".join(f"{k}={v}" for k, v in p.items()) or "default")
def test_bands_equal_whole_frame_all_policies(policy, cases):
    env = dict(policy, SR_POISON="1")
    os.environ.update(env)
    try:
        r = srb.CudaRenderer(0)
        for case, (v, i, m, W, H, bands, want) in enumerate(cases):
            for dt in (False, True):
                c, d = render_in_bands(r, v, i, m, W, H, bands, depth_test=dt)
                rc_c, rc_d = want[dt]
                bad = (c != rc_c).any(-1)
                assert not bad.any(), f"case {case} {W}x{H} bands {bands} depth={dt}: {int(bad.sum())} colour pixels differ, first {np.argwhere(bad)[0]}"
                assert np.array_equal(d.view(np.uint32), rc_d.view(np.uint32)), f"case {case} bands {bands} depth={dt}: depth differs"
        r.close()
    finally:
        for k in env:
            del os.environ[k]


def test_bands_far_offscreen_corners(renderer):
    """Corners up to 2^29 pixels away: Float(Int) of the frame row is inexact there, so shading must use frame rows,
    not band rows (the band shift is applied to the integer span rule only)."""
    rng = np.random.default_rng(77)
    n = 40
    xyz = rng.uniform(-1.0, 1.0, size=(3 * n, 3))
    far = rng.random(3 * n) < 0.3
    xyz[far, :2] *= np.exp(rng.uniform(np.log(1e2), np.log(1.5e6), size=(int(far.sum()), 1)))
    xyz[:, 2] = rng.uniform(0.0, 1.0, size=3 * n)
    v = S.pack_vertices(xyz.astype(np.float32), rng.uniform(0, 1, size=(3 * n, 3)).astype(np.float32))
    i = np.arange(3 * n, dtype=np.int64)
    m = np.eye(4, dtype=np.float32)
    W, H = 300, 200
    for dt in (False, True):
        rc, rc_c, rc_d, _ = O.render(v, i, m, W, H, flags=O.DEPTH_LESS if dt else 0)
        assert rc == O.OK
        c, d = render_in_bands(renderer, v, i, m, W, H, [(0, 7), (7, 100), (100, 101), (101, 200)], depth_test=dt)
        assert np.array_equal(c, rc_c) and np.array_equal(d.view(np.uint32), rc_d.view(np.uint32))


@pytest.mark.parametrize("prim", [srb.PrimitiveType.vertices, srb.PrimitiveType.line])
def test_bands_points_and_lines(renderer, prim):
    rng = np.random.default_rng(5)
    n = 3000
    xyz = rng.uniform(-1.1, 1.1, size=(n, 3)).astype(np.float32)
    v = S.pack_vertices(xyz, rng.uniform(0, 1, size=(n, 3)).astype(np.float32))
    i = rng.integers(0, n, size=3 * 2000).astype(np.int64)
    W, H = 211, 157
    rc, rc_c, rc_d, _ = O.render(v, i, np.eye(4, dtype=np.float32), W, H, primitive_type=int(prim))
    assert rc == O.OK
    c, d = render_in_bands(renderer, v, i, np.eye(4, dtype=np.float32), W, H, [(0, 50), (50, 51), (51, 157)], prim=prim)
    assert np.array_equal(c, rc_c) and np.array_equal(d.view(np.uint32), rc_d.view(np.uint32))


def test_band_sized_outputs_of_the_batch_and_device_calls(renderer):
    """render_views / render_device with a band set write images of the band's rows only."""
    import torch
    v, i = S.icosphere(4)
    views = np.stack([S.app_transform(t) for t in (0.1, 0.9, 2.3)])
    W, H = 320, 200
    renderer.set_mesh(v, i)
    full_c, full_d = renderer.render_views(views, W, H)
    for world in (2, 3, 8):
        parts_c, parts_d = [], []
        for rank in range(world):
            y0, y1 = band_range(H, rank, world)
            if y0 == y1:
                continue
            renderer.set_band(y0, y1)
            c, d = renderer.render_views(views, W, H)
            assert c.shape == (3, y1 - y0, W, 4) and d.shape == (3, y1 - y0, W)
            dc = torch.full((3, y1 - y0, W, 4), 9, dtype=torch.uint8, device="cuda:0")
            dd = torch.full((3, y1 - y0, W), 9.0, dtype=torch.float32, device="cuda:0")
            torch.cuda.synchronize()                 # the library renders on its own stream
            renderer.render_device(views, W, H, dc.data_ptr(), dd.data_ptr())
            assert np.array_equal(dc.cpu().numpy(), c) and np.array_equal(dd.cpu().numpy().view(np.uint32), d.view(np.uint32))
            parts_c.append(c)
            parts_d.append(d)
        renderer.set_band()
        assert np.array_equal(np.concatenate(parts_c, axis=1), full_c)
        assert np.array_equal(np.concatenate(parts_d, axis=1).view(np.uint32), full_d.view(np.uint32))


def test_soup_frame_in_bands_equals_whole_frame(renderer):
    """BASELINE-size property: the 1M-triangle soup at 1080p rendered as 8 bands (batched frames, tile retirement and
    slab growth all happen per band) stitches to the whole frame, depth off and on."""
    W, H = 1920, 1080
    v, i = S.triangle_soup(1_000_000, W, H)
    m = S.soup_transform(3)
    renderer.set_mesh(v, i)
    for dt in (False, True):
        renderer.depth_test = dt
        try:
            full_c, full_d = renderer.render_views(m[None], W, H)
            full_c, full_d = renderer.render_views(m[None], W, H)       # second frame: the adaptive policies have settled
            rows_c, rows_d = [], []
            for rank in range(8):
                renderer.set_band(*band_range(H, rank, 8))
                for _ in range(2):
                    c, d = renderer.render_views(m[None], W, H)
                rows_c.append(c)
                rows_d.append(d)
        finally:
            renderer.set_band()
            renderer.depth_test = False
        assert np.array_equal(np.concatenate(rows_c, axis=1), full_c), f"depth={dt}"
        assert np.array_equal(np.concatenate(rows_d, axis=1).view(np.uint32), full_d.view(np.uint32)), f"depth={dt}"


def test_band_argument_errors(renderer):
    v, i = S.icosphere(1)
    renderer.set_mesh(v, i)
    with pytest.raises(srb.SrError) as e:
        renderer.set_band(5, 3)
    assert e.value.code == srb.renderer.SR_ERR_ARG
    with pytest.raises(srb.SrError):
        renderer.set_band(-1, 3)
    renderer.set_band(10, 300)
    try:
        with pytest.raises(srb.SrError) as e:
            renderer.render_views(np.eye(4, dtype=np.float32)[None], 64, 200)     # the band ends below a 200-row image
        assert e.value.code == srb.renderer.SR_ERR_ARG
    finally:
        renderer.set_band()
    c, _ = renderer.render_views(np.eye(4, dtype=np.float32)[None], 64, 200)
    assert c.shape == (1, 200, 64, 4)
