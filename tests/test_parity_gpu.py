"""GPU parity tests: the CUDA path, called through the C ABI (ctypes), against the CPU oracle on the same inputs.
Bar: coverage, depth and colour all bit-exact (integer/byte work; the float arithmetic is the same sequence of
individually rounded operations on both sides)."""
import numpy as np
import pytest

import software_renderer_b200 as srb
from oracle import oracle as O
from software_renderer_b200 import scene as S

pytestmark = pytest.mark.gpu
I4 = np.eye(4, dtype=np.float32)


def gpu_render(r, v, i, m, W, H, prim=srb.PrimitiveType.triangle, depth_test=False, expect_domain=False):
    r.depth_test = depth_test
    rp = srb.RenderPass(srb.Image.color(W, H), srb.Image.depth(W, H), v, i, prim, m)
    rp.color_buffer.array[:] = 0x5A      # stale contents must be overwritten (clear is part of the pass)
    rp.depth_buffer.array[:] = -7.0
    if expect_domain:
        with pytest.raises(srb.DomainError):
            r.render(rp)
    else:
        r.render(rp)
    r.depth_test = False
    return rp.color_buffer.array, rp.depth_buffer.array


def assert_same(r, v, i, m, W, H, prim=srb.PrimitiveType.triangle, depth_test=False, expect_domain=False):
    c, d = gpu_render(r, v, i, m, W, H, prim, depth_test, expect_domain)
    flags = O.DEPTH_LESS if depth_test else 0
    rc, rc_c, rc_d, st = O.render(v, i, m, W, H, primitive_type=int(prim), flags=flags)
    assert rc == (O.ERR_DOMAIN if expect_domain else O.OK)
    bad = (c != rc_c).any(-1)
    assert not bad.any(), f"{int(bad.sum())} colour pixels differ, first at {np.argwhere(bad)[0]}"
    assert np.array_equal(d.view(np.uint32), rc_d.view(np.uint32)), "depth differs"
    return c, d, st


def tri(xy):
    xyz = np.array([[x, y, 0.0] for x, y in xy], np.float32)
    return S.pack_vertices(xyz, np.eye(3, dtype=np.float32)), np.array([0, 1, 2], np.int64)


def random_tris(seed, n, spread=1.3, zspread=(0.0, 1.0), small=None):
    rng = np.random.default_rng(seed)
    if small is None:
        xyz = rng.uniform(-spread, spread, size=(3 * n, 3))
    else:
        c = rng.uniform(-spread, spread, size=(n, 1, 3))
        xyz = (c + rng.uniform(-small, small, size=(n, 3, 3))).reshape(-1, 3)
    xyz[:, 2] = rng.uniform(zspread[0], zspread[1], size=3 * n)
    rgb = rng.uniform(-0.2, 1.2, size=(3 * n, 3))
    return S.pack_vertices(xyz.astype(np.float32), rgb.astype(np.float32)), np.arange(3 * n, dtype=np.int64)


def test_kat1_and_kat2(renderer):
    v, i = tri([(0, 1), (1, 0), (-1, 0)])
    c, d, st = assert_same(renderer, v, i, I4, 512, 512)
    assert (c[..., 3] > 0).sum() == 65537
    assert c[0, 256].tolist() == [0, 0, 255, 255] and c[128, 256].tolist() == [63, 63, 127, 255]
    assert c[256, 0].tolist() == [255, 0, 0, 255] and c[100, 300].tolist() == [27, 71, 155, 255]
    assert np.isposinf(d).all()
    v, i = tri([(0, 0.5), (-0.5, -0.5), (0.5, -0.5)])
    c, d, st = assert_same(renderer, v, i, I4, 512, 512)
    assert np.nonzero(c[384, :, 3])[0].tolist() == [384]


@pytest.mark.parametrize("depth_test", [False, True])
@pytest.mark.parametrize("seed,n,size,kw", [
    (1, 60, (97, 61), {}),                       # big triangles, odd image size (partial tiles)
    (2, 400, (256, 256), {"small": 0.03}),       # small triangles: direct path
    (3, 300, (640, 360), {"small": 0.15}),       # mixed
    (4, 200, (333, 777), {"spread": 3.0}),       # mostly off-screen / partially clipped
    (5, 2000, (800, 600), {"small": 0.004}),     # sub-pixel triangles (integer-degenerate -> NaN weights)
    (6, 50, (1920, 1080), {}),                   # full-screen triangles, heavy overdraw
])
def test_random_triangles(renderer, seed, n, size, kw, depth_test):
    v, i = random_tris(seed, n, **kw)
    assert_same(renderer, v, i, I4, size[0], size[1], depth_test=depth_test)


@pytest.mark.parametrize("depth_test", [False, True])
def test_degenerate_and_tie_cases(renderer, depth_test):
    # equal-y rows, coincident vertices, vertical slivers, exact depth ties between overlapping triangles
    xy = [(-0.5, 0.0), (0.9, 0.0), (0.0, 0.0),   (0.2, 0.2), (0.2, 0.2), (0.2, 0.2),   (0.5, -0.9), (0.5, 0.9), (0.5, 0.1),
          (-1, -1), (1, -1), (0, 1),   (-1, -1), (1, -1), (0, 1),   (1, 1), (-1, 1), (0, -1)]
    xyz = np.array([[x, y, 0.5] for x, y in xy], np.float32)
    rng = np.random.default_rng(0)
    v = S.pack_vertices(xyz, rng.uniform(0, 1, size=(len(xy), 3)).astype(np.float32))
    i = np.arange(len(xy), dtype=np.int64)
    assert_same(renderer, v, i, I4, 200, 120, depth_test=depth_test)
    assert_same(renderer, v, i[::-1].copy(), I4, 200, 120, depth_test=depth_test)


def test_order_defines_visibility(renderer):
    v, i = random_tris(11, 30)
    c1, _, _ = assert_same(renderer, v, i, I4, 320, 200)
    j = i.reshape(-1, 3)[::-1].reshape(-1).copy()
    c2, _, _ = assert_same(renderer, v, j, I4, 320, 200)
    assert not np.array_equal(c1, c2)


@pytest.mark.parametrize("k", [0, 13, 47])
def test_config1_cube_800x600(renderer, k):
    v, i = S.cube()
    assert_same(renderer, v, i, S.app_transform(k / 60), 800, 600)
    assert_same(renderer, v, i, S.app_transform(k / 60), 800, 600, depth_test=True)


@pytest.mark.parametrize("depth_test", [False, True])
def test_config2_icosphere_1080p(renderer, depth_test):
    v, i = S.icosphere(6)
    c, d, st = assert_same(renderer, v, i, S.app_transform(0.5), 1920, 1080, depth_test=depth_test)
    assert st["primitives"] == 81920


@pytest.mark.parametrize("depth_test", [False, True])
def test_config3_soup_reduced(renderer, depth_test):
    v, i = S.triangle_soup(60_000, 1920, 1080, seed=3)
    assert_same(renderer, v, i, S.soup_transform(2), 1920, 1080, depth_test=depth_test)


def test_config3_soup_full_1m(renderer):
    v, i = S.triangle_soup(1_000_000, 1920, 1080)
    c, d, st = assert_same(renderer, v, i, S.soup_transform(0), 1920, 1080)
    assert (c[..., 3] > 0).all()
    s = renderer.stats()
    assert s["small_tris"] + s["binned_tris"] <= 1_000_000 and s["binned_tris"] > 100_000


def test_config3_soup_full_1m_depth_less(renderer):
    """BASELINE.json calls configs[2] the "depth-test stress": the full 1M-triangle soup with the test of
    Renderer.swift:257-261 on, two consecutive frames (the second runs with the launch policies the first one set)."""
    v, i = S.triangle_soup(1_000_000, 1920, 1080)
    for k in (0, 1):
        c, d, st = assert_same(renderer, v, i, S.soup_transform(k), 1920, 1080, depth_test=True)
    assert np.isfinite(d[c[..., 3] > 0]).all()
    assert renderer.stats()["depth_culled"] > 0


@pytest.mark.parametrize("depth_test", [False, True])
def test_config5_cubesphere_512_views(renderer, depth_test):
    """configs[4]: 32 of the 4096 camera views, spread over the whole list, in one batched call."""
    v, i = S.cube_sphere(289)
    picks = sorted(set(int(x) for x in np.linspace(0, 4095, 32)))
    views = S.fibonacci_views(4096)[picks]
    renderer.set_mesh(v, i)
    renderer.depth_test = depth_test
    try:
        color, depth = renderer.render_views(views, 512, 512)
    finally:
        renderer.depth_test = False
    for k in range(len(picks)):
        rc, rc_c, rc_d, _ = O.render(v, i, views[k], 512, 512, flags=O.DEPTH_LESS if depth_test else 0)
        assert rc == 0 and np.array_equal(color[k], rc_c), f"view {picks[k]}"
        assert np.array_equal(depth[k].view(np.uint32), rc_d.view(np.uint32)), f"view {picks[k]}"


def test_config4_cubesphere_10m_4k_properties(renderer):
    """Full size of configs[3]: the oracle needs ~10 s here, so one frame is compared directly."""
    v, i = S.cube_sphere(913)
    assert i.size // 3 == 10_002_828
    m = S.app_transform(0.25)
    assert_same(renderer, v, i, m, 3840, 2160)


def test_vertices_and_line_primitives(renderer):
    v, i = random_tris(21, 500, spread=1.2)
    assert_same(renderer, v, i, I4, 300, 200, prim=srb.PrimitiveType.vertices)
    c, d, _ = assert_same(renderer, v, i, I4, 300, 200, prim=srb.PrimitiveType.line)
    assert not c.any()
    # collisions: many corners on one pixel -> the last corner wins
    xyz = np.zeros((30, 3), np.float32)
    rgb = np.random.default_rng(5).uniform(0, 1, size=(30, 3)).astype(np.float32)
    assert_same(renderer, S.pack_vertices(xyz, rgb), np.arange(30, dtype=np.int64), I4, 64, 64, prim=srb.PrimitiveType.vertices)


def test_domain_rule_and_index_errors(renderer):
    v, i = random_tris(31, 40)
    v[5, 0] = np.inf
    v[40, 1] = np.nan
    v[77, 0] = 3e9           # |sx| >= 2^30
    assert_same(renderer, v, i, I4, 128, 128, expect_domain=True)
    assert renderer.stats()["domain_skipped"] == 3
    assert_same(renderer, v, i, I4, 128, 128, prim=srb.PrimitiveType.vertices, expect_domain=True)
    m = I4.copy()
    m[3, 3] = 0.0            # w = 0 for every vertex
    assert_same(renderer, *random_tris(32, 10), m, 64, 64, expect_domain=True)
    bad = i.copy()
    bad[10] = v.shape[0]
    with pytest.raises(srb.SrError) as e:
        gpu_render(renderer, v, bad, I4, 64, 64)
    assert e.value.code == -2
    ok_trailing = np.concatenate([np.arange(6), [10 ** 9]]).astype(np.int64)   # partial primitive is ignored
    assert_same(renderer, *random_tris(33, 2)[:1], ok_trailing, I4, 64, 64)


def test_empty_mesh_is_a_clear(renderer):
    v = np.zeros((0, 8), np.float32)
    i = np.zeros((0,), np.int64)
    c, d, _ = assert_same(renderer, v, i, I4, 50, 40)
    assert not c.any() and np.isposinf(d).all()


def test_batch_equals_single_and_device_path(renderer):
    import torch
    v, i = S.icosphere(4)
    views = S.fibonacci_views(16)
    renderer.set_mesh(v, i)
    color, depth = renderer.render_views(views, 256, 192)
    dev_c = torch.empty((16, 192, 256, 4), dtype=torch.uint8, device="cuda")
    dev_d = torch.empty((16, 192, 256), dtype=torch.float32, device="cuda")
    renderer.render_device(views, 256, 192, dev_c.data_ptr(), dev_d.data_ptr(), stream=torch.cuda.current_stream().cuda_stream)
    assert np.array_equal(dev_c.cpu().numpy(), color) and np.array_equal(dev_d.cpu().numpy(), depth)
    for k in (0, 7, 15):
        rc, rc_c, rc_d, _ = O.render(v, i, views[k], 256, 192)
        assert np.array_equal(color[k], rc_c) and np.array_equal(depth[k], rc_d)


@pytest.mark.parametrize("env", [{"SR_SLAB": "0"}, {"SR_SLAB_CAP": "4"}, {}])
def test_pair_buffer_regrow(env):
    """Many full-screen triangles overflow the initial tile lists (compact lists: the pair buffer; slab lists: the
    per-(tile, chunk) segments); the frame is re-run and still exact."""
    import os
    os.environ.update(env)
    try:
        r = srb.CudaRenderer(0)
    finally:
        for k in env:
            del os.environ[k]
    try:
        v, i = random_tris(41, 1000, spread=1.5)
        for dt in (False, True):
            c, d = gpu_render(r, v, i, I4, 1920, 1080, depth_test=dt)
            rc, rc_c, rc_d, _ = O.render(v, i, I4, 1920, 1080, flags=O.DEPTH_LESS if dt else 0)
            assert np.array_equal(c, rc_c) and np.array_equal(d.view(np.uint32), rc_d.view(np.uint32))
            if not dt:
                assert r.stats()["tile_pairs"] > (1 << 20)
                assert r.stats()["bin_regrows"] >= 1 or not env
    finally:
        r.close()


@pytest.mark.parametrize("depth_test", [False, True])
def test_compact_lists_mode_is_identical(depth_test):
    """SR_SLAB=0: the scan + fill binning path (the fallback when slabs would exceed their memory limit)."""
    import os
    os.environ["SR_SLAB"] = "0"
    try:
        r = srb.CudaRenderer(0)
    finally:
        del os.environ["SR_SLAB"]
    try:
        v, i = S.triangle_soup(120_000, 1280, 720, seed=21)
        for k in range(3):
            assert_same(r, v, i, S.soup_transform(k), 1280, 720, depth_test=depth_test)
        v, i = S.icosphere(5)
        assert_same(r, v, i, S.app_transform(0.2), 1000, 700, depth_test=depth_test)
    finally:
        r.close()


def test_repeatability_and_self_clearing(renderer):
    v, i = random_tris(51, 200, small=0.2)
    a = [gpu_render(renderer, v, i, S.app_transform(t), 400, 300, depth_test=dt) for t in (0.1, 0.2) for dt in (False, True)]
    b = [gpu_render(renderer, v, i, S.app_transform(t), 400, 300, depth_test=dt) for t in (0.1, 0.2) for dt in (False, True)]
    for (c1, d1), (c2, d2) in zip(a, b):
        assert np.array_equal(c1, c2) and np.array_equal(d1, d2)


@pytest.fixture(scope="module")
def batched_renderer():
    """A context forced into index-descending batch mode (SR_BATCHES=1), which the adaptive policy otherwise only
    picks for heavy-overdraw frames."""
    import os
    os.environ["SR_BATCHES"] = "1"
    try:
        r = srb.CudaRenderer(0)
    finally:
        del os.environ["SR_BATCHES"]
    yield r
    r.close()


@pytest.mark.parametrize("case", ["soup", "soup_rot", "ico", "big", "sparse", "cube_sphere"])
def test_batched_mode_is_identical(batched_renderer, case):
    r = batched_renderer
    if case == "soup":
        v, i = S.triangle_soup(200_000, 1280, 720, seed=9)
        assert_same(r, v, i, S.soup_transform(1), 1280, 720)
        assert r.stats()["batches_run"] >= 1
    elif case == "soup_rot":
        v, i = S.triangle_soup(150_000, 801, 603, seed=10)      # partial tiles + two consecutive frames
        assert_same(r, v, i, S.soup_transform(0), 801, 603)
        assert_same(r, v, i, S.soup_transform(40), 801, 603)
    elif case == "ico":
        v, i = S.icosphere(6)
        assert_same(r, v, i, S.app_transform(0.3), 1920, 1080)
    elif case == "big":
        v, i = random_tris(61, 70_000, spread=1.2, small=0.3)   # large triangles, every tile retires early
        assert_same(r, v, i, I4, 640, 480)
        s = r.stats()
        assert s["batches_run"] < 2 and s["pairs_consumed"] < s["tile_pairs"]
    elif case == "sparse":
        v, i = random_tris(62, 70_000, spread=0.4, small=0.01)  # most of the screen stays background
        assert_same(r, v, i, I4, 640, 480)
        assert r.stats()["batches_run"] == 2
    else:
        v, i = S.cube_sphere(120)
        assert_same(r, v, i, S.app_transform(0.7), 1024, 1024)


@pytest.mark.parametrize("W", [4080, 4096])
def test_batched_wide_frames_at_the_box_word_limit(batched_renderer, W):
    """A box word of the late-batch pre-pass (k_late_boxes) holds tile coordinates up to 254: a frame of 255 tile columns
    (4080 pixels) still takes it, one of 256 the path that transforms every late triangle itself.  Both are the oracle's image."""
    v, i = random_tris(71, 70_000, spread=1.1, small=0.02)
    assert_same(batched_renderer, v, i, I4, W, 48)
    assert batched_renderer.stats()["batches_run"] >= 1


@pytest.mark.parametrize("case", ["sparse", "all_retired"])
def test_batched_domain_skips_of_the_later_batch_are_counted(batched_renderer, case):
    """Triangles outside the valid domain (the reference traps, R:251 / R:271) among the LOW indices, i.e. in the later batch of
    a batched frame: skipped, counted once each, the call reports the domain error - also when the first batch retires
    every tile and the later batch's own kernels leave at once (the count then comes from the box pre-pass)."""
    if case == "sparse":
        v, i = random_tris(72, 70_000, spread=0.4, small=0.01)
    else:
        v, i = random_tris(73, 70_000, spread=1.2, small=0.3)
    v = v.copy()
    for t in (5, 6, 7):
        v[3 * t + 1, 0] = np.inf                      # one corner of triangles 5, 6, 7 has no finite screen position
    assert_same(batched_renderer, v, i, I4, 640, 480, expect_domain=True)
    assert batched_renderer.stats()["domain_skipped"] == 3
    assert batched_renderer.stats()["batches_run"] == (2 if case == "sparse" else 1)


def test_adaptive_policy_switches_and_stays_exact(renderer):
    v, i = S.triangle_soup(300_000, 1280, 720, seed=12)
    seen = set()
    for k in range(4):
        assert_same(renderer, v, i, S.soup_transform(k), 1280, 720)
        seen.add(renderer.stats()["batches_run"] > 0)
    assert seen == {False, True}            # first frame single-batch, later frames batched
    v, i = S.icosphere(5)
    for k in range(3):
        assert_same(renderer, v, i, S.app_transform(0.1 * k), 1280, 720)


@pytest.mark.parametrize("depth_test", [False, True])
def test_nobin_speculation_and_rerun(depth_test):
    """After a frame without large triangles the binning kernels are not launched; a later frame that does meet a
    large triangle is re-run with the full pipeline.  Every frame stays bit-exact."""
    r = srb.CudaRenderer(0)
    try:
        v, i = S.cube_sphere(40)                     # 19,200 triangles
        zoom = lambda s: (np.diag([s, s, 1.0, 1.0]).astype(np.float32) @ S.app_transform(0.3)).astype(np.float32)
        W, H = 640, 640
        assert_same(r, v, i, zoom(0.25), W, H, depth_test=depth_test)          # everything small: full pipeline, then speculate
        assert r.stats()["binned_tris"] == 0
        n_full = r.stats()["gpu_launches"]
        assert_same(r, v, i, zoom(0.26), W, H, depth_test=depth_test)          # speculative frame: fewer launches
        assert r.stats()["gpu_launches"] <= n_full and r.stats()["nobin_reruns"] == 0
        assert_same(r, v, i, zoom(1.5), W, H, depth_test=depth_test)           # large triangles appear: re-run
        assert r.stats()["nobin_reruns"] == 1 and r.stats()["binned_tris"] > 0
        assert_same(r, v, i, zoom(1.5), W, H, depth_test=depth_test)           # policy went back to the full pipeline
        assert r.stats()["nobin_reruns"] == 0
        views = np.stack([zoom(s) for s in (0.2, 0.25, 1.2, 0.22)])            # a batch in which one view fails
        assert_same(r, v, i, zoom(0.2), W, H, depth_test=depth_test)           # all small again -> speculate next call
        r.depth_test = depth_test
        color, depth = r.render_views(views, W, H)
        assert r.stats()["nobin_reruns"] == 1
        r.depth_test = False
        for k in range(4):
            rc, rc_c, rc_d, _ = O.render(v, i, views[k], W, H, flags=O.DEPTH_LESS if depth_test else 0)
            assert np.array_equal(color[k], rc_c) and np.array_equal(depth[k].view(np.uint32), rc_d.view(np.uint32)), k
    finally:
        r.close()


def test_tiny_triangle_path_all_shapes(renderer):
    """Every triangle whose integer corners fit a 2x2 pixel box (the lane-local path of k_setup), in every corner
    order, plus neighbours that are one pixel wider or taller (the warp-balanced path), both depth modes."""
    pts = [(x, y) for x in range(3) for y in range(3)]
    tris = [(a, b, c) for a in pts for b in pts for c in pts]
    n = len(tris)
    W = H = 4 * 28
    xyz = np.zeros((3 * n, 3), np.float32)
    rng = np.random.default_rng(5)
    for k, t in enumerate(tris):
        ox, oy = 4 * (k % 28) + 0.3, 4 * (k // 28) + 0.4
        for q, (px, py) in enumerate(t):
            sx, sy = ox + px + 0.2 * q, oy + py + 0.1 * q          # fractional screen coordinates, distinct float y
            xyz[3 * k + q] = ((sx / W - 0.5) / 0.5, (sy / H - 0.5) / -0.5, rng.uniform(0.1, 0.9))
    v = S.pack_vertices(xyz, rng.uniform(0, 1, size=(3 * n, 3)).astype(np.float32))
    i = np.arange(3 * n, dtype=np.int64)
    for dt in (False, True):
        assert_same(renderer, v, i, I4, W, H, depth_test=dt)
    # the same shapes hanging over the image border
    shift = np.eye(4, dtype=np.float32); shift[0, 3] = -1.0 + 1.0 / W; shift[1, 3] = 1.0 - 1.0 / H
    assert_same(renderer, v, i, shift, W, H)


@pytest.mark.parametrize("size", [(641, 479), (33, 7), (2051, 2049), (2048, 2050)])
@pytest.mark.parametrize("depth_test", [False, True])
def test_resolve_kernel_odd_sizes(size, depth_test):
    """The no-binning resolve kernel handles four pixels per thread: widths and pixel counts that are not multiples
    of four, rows straddled by one thread's four pixels."""
    W, H = size
    r = srb.CudaRenderer(0)
    try:
        v, i = S.cube_sphere(60 if W < 1000 else 420)
        m = (np.diag([0.5, 0.5, 1.0, 1.0]).astype(np.float32) @ S.app_transform(0.9)).astype(np.float32)
        for k in range(3):                                # frame 0 full pipeline, frames 1.. speculative (k_resolve)
            assert_same(r, v, i, m, W, H, depth_test=depth_test)
            assert r.stats()["nobin_reruns"] == 0 and r.stats()["binned_tris"] == 0
    finally:
        r.close()


def _layered(n, seed, zfun, spread=1.2, size=0.9):
    rng = np.random.default_rng(seed)
    c = rng.uniform(-spread, spread, size=(n, 1, 2))
    xy = (c + rng.uniform(-size, size, size=(n, 3, 2))).reshape(-1, 2)
    z = zfun(rng, n).astype(np.float64).reshape(-1)
    xyz = np.concatenate([xy, z[:, None]], 1).astype(np.float32)
    rgb = rng.uniform(0, 1, size=(3 * n, 3)).astype(np.float32)
    return S.pack_vertices(xyz, rgb), np.arange(3 * n, dtype=np.int64)


@pytest.mark.parametrize("case", ["front_last", "front_first", "ties", "tilted", "slivers", "huge_z", "negative_w"])
def test_depth_bounds_never_change_the_image(case):
    """Depth mode rejects (triangle, tile) records by a conservative depth bound and stops a tile's front-to-back
    list early; the image must stay bit-exact whatever the layering, including exact ties, slivers with huge
    barycentric weights and non-finite depths."""
    n = 3000
    W, H = 640, 480
    m = I4
    if case == "front_last":            # nearest triangles have the highest indices
        v, i = _layered(n, 1, lambda r, k: np.repeat(np.linspace(0.9, 0.1, k), 3))
    elif case == "front_first":
        v, i = _layered(n, 2, lambda r, k: np.repeat(np.linspace(0.1, 0.9, k), 3))
    elif case == "ties":                # a few exact depth levels: the earliest index must win every tie
        v, i = _layered(n, 3, lambda r, k: np.repeat(r.integers(0, 4, size=k) * 0.25, 3))
    elif case == "tilted":              # per-corner depths: planes intersect inside tiles
        v, i = _layered(n, 4, lambda r, k: r.uniform(0.0, 1.0, size=3 * k))
    elif case == "slivers":
        v, i = _layered(n, 5, lambda r, k: r.uniform(-1.0, 1.0, size=3 * k))
        xyz = v[:, :3].copy().reshape(-1, 3, 3)
        xyz[:, 2, :2] = xyz[:, 1, :2] + (xyz[:, 1, :2] - xyz[:, 0, :2]) * 1e-3      # nearly collinear corners
        v = S.pack_vertices(xyz.reshape(-1, 3), v[:, 4:7])
    elif case == "huge_z":              # products overflow to inf, inf - inf gives NaN depths: neither ever passes R:258
        def z(r, k):
            out = r.uniform(0.0, 1.0, size=3 * k)
            out[r.integers(0, 3 * k, size=k // 8)] = 3.0e38
            out[r.integers(0, 3 * k, size=k // 8)] = -3.0e38
            out[r.integers(0, 3 * k, size=k // 8)] = 1.0e-42
            return out
        v, i = _layered(n, 6, z)
    else:                               # perspective with w = z + 1 crossing zero
        v, i = _layered(n, 7, lambda r, k: r.uniform(-1.6, 1.0, size=3 * k), spread=0.8, size=0.5)
        m = np.array([[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, 1, 0], [0, 0, 1, 1]], np.float32)
    r = srb.CudaRenderer(0)
    try:
        for k in range(2):              # second frame uses thresholds learned from the first
            with np.errstate(all="ignore"):
                assert_same(r, v, i, m, W, H, depth_test=True)
        if case == "front_first":
            s = r.stats()
            assert s["depth_culled"] > 0.5 * s["tile_pairs"]
    finally:
        r.close()


def test_compact_lists_right_edge_columns():
    """Compact tile lists: the counting pass (k_setup) and the fill pass (k_bin_fill) must agree tile by tile, or a
    list has a gap of never-written entries.  Triangles whose rows leave the image through a partial last tile
    column (width not a multiple of 16) are the case where the two passes once disagreed."""
    import os
    env = {"SR_SLAB": "0", "SR_POISON": "1"}
    os.environ.update(env)
    try:
        r = srb.CudaRenderer(0)
        W, H = 690, 400
        rng = np.random.default_rng(5)
        tris = []
        for _ in range(200):
            x0 = rng.uniform(660, 689); y0 = rng.uniform(0, 200)
            tris += [(x0, y0), (x0 + rng.uniform(100, 400), y0 + rng.uniform(0, 30)), (x0 + rng.uniform(100, 400), y0 + rng.uniform(100, 300))]
        px = np.array(tris)
        xyz = np.stack([px[:, 0] / W * 2 - 1, 1 - px[:, 1] / H * 2, rng.uniform(0, 1, len(px))], 1).astype(np.float32)
        v = S.pack_vertices(xyz, rng.uniform(0, 1, size=(len(px), 3)).astype(np.float32))
        i = np.arange(len(px), dtype=np.int64)
        for dt in (False, True):
            assert_same(r, v, i, I4, W, H, depth_test=dt)
        assert r.stats()["binned_tris"] > 0
        r.close()
    finally:
        for k in env:
            del os.environ[k]
