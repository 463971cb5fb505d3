"""Round-2 boundary behaviour through the C ABI: the strict per-call mesh hand-over (App.swift:163), the static-mesh
opt-in, page-locked and pageable caller memory, the depth plane of passes that never test depth (Renderer.swift:206),
size limits, alignment of device outputs, frame-epoch wrap."""
import ctypes
import os
import subprocess
import sys

import numpy as np
import pytest

import software_renderer_b200 as srb
from oracle import oracle as O
from software_renderer_b200 import scene as S

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


def oracle_images(v, i, m, W, H, depth=False):
    rc, c, d, _ = O.render(v, i, m, W, H, flags=O.DEPTH_LESS if depth else 0)
    assert rc == O.OK
    return c, d


def same(rp, v, i, m, W, H, depth=False):
    c, d = oracle_images(v, i, m, W, H, depth)
    return np.array_equal(rp.color_buffer.array, c) and np.array_equal(rp.depth_buffer.array.view(np.uint32), d.view(np.uint32))


def test_default_pass_uploads_the_mesh_every_call_so_in_place_edits_are_seen(renderer):
    v, i = S.icosphere(3)
    m = S.app_transform(0.3)
    W, H = 400, 300
    rp = srb.RenderPass(srb.Image.color(W, H), srb.Image.depth(W, H), v, i, srb.PrimitiveType.triangle, m)
    renderer.render(rp)
    assert same(rp, v, i, m, W, H)
    g0 = renderer._lib.sr_mesh_generation(renderer._ctx)
    v[:, 0] *= 0.5                       # same array object, same address, same size: only the contents change
    v[:, 4:7] = 1.0 - v[:, 4:7]
    renderer.render(rp)
    assert same(rp, v, i, m, W, H)
    assert renderer._lib.sr_mesh_generation(renderer._ctx) == g0 + 1


def test_static_mesh_opt_in_uploads_on_a_version_bump_only(renderer):
    v, i = S.icosphere(3)
    m = S.app_transform(0.3)
    W, H = 400, 300
    rp = srb.RenderPass(srb.Image.color(W, H), srb.Image.depth(W, H), v, i, srb.PrimitiveType.triangle, m, mesh_version=1)
    renderer.render(rp)
    g0 = renderer._lib.sr_mesh_generation(renderer._ctx)
    before = rp.color_buffer.array.copy()
    renderer.render(rp)
    assert renderer._lib.sr_mesh_generation(renderer._ctx) == g0          # no upload
    old_v = v.copy()
    v[:, 0] *= 0.5
    renderer.render(rp)                                                     # the caller did not bump: the old mesh
    assert np.array_equal(rp.color_buffer.array, before) and same(rp, old_v, i, m, W, H)
    rp.mesh_version = 2
    renderer.render(rp)
    assert same(rp, v, i, m, W, H) and renderer._lib.sr_mesh_generation(renderer._ctx) == g0 + 1
    # another renderer call that replaces the mesh invalidates the cached key (library generation)
    renderer.set_mesh(*S.cube())
    renderer.render(rp)
    assert same(rp, v, i, m, W, H)


def test_page_locked_and_pageable_images_and_meshes(renderer):
    v, i = S.triangle_soup(60_000, 640, 360, seed=9)          # 5.8 MB of vertices: the bounce-buffer path for pageable arrays
    m = S.soup_transform(1)
    W, H = 640, 360
    c, d = oracle_images(v, i, m, W, H)
    pv = renderer.host_array(v.shape, np.float32)
    pi = renderer.host_array(i.shape, np.int64)
    pv[:] = v
    pi[:] = i
    pc = renderer.host_array((H, W, 4), np.uint8)
    pd = renderer.host_array((H, W), np.float32)
    for verts, idx, col, dep in ((v, i, np.empty((H, W, 4), np.uint8), np.empty((H, W), np.float32)), (pv, pi, pc, pd)):
        col[:] = 0x33
        dep[:] = -1.0
        rp = srb.RenderPass(srb.Image(col, W, H), srb.Image(dep, W, H), verts, idx, srb.PrimitiveType.triangle, m)
        renderer.render(rp)
        assert np.array_equal(col, c) and np.array_equal(dep.view(np.uint32), d.view(np.uint32))
    big = np.zeros(3 << 20, np.uint8)
    renderer.host_register(big)
    renderer.host_unregister(big)


def test_depth_plane_of_untested_passes(renderer):
    """Depth test off: the pass leaves +inf everywhere (R:206).  Default: written on every call (by the helper threads,
    not over PCIe).  SR_FLAG_DEPTH_UNTOUCHED: not written again while the caller keeps handing over the same image."""
    v, i = S.icosphere(3)
    W, H = 320, 240
    col, dep = np.empty((H, W, 4), np.uint8), np.full((H, W), -3.0, np.float32)
    rp = srb.RenderPass(srb.Image(col, W, H), srb.Image(dep, W, H), v, i, srb.PrimitiveType.triangle, S.app_transform(0.1), mesh_version=7)
    renderer.render(rp)
    assert np.isposinf(dep).all()
    dep[5, 5] = 2.0
    renderer.render(rp)
    assert np.isposinf(dep).all()                          # default: rewritten every call
    try:
        renderer.depth_untouched = True
        renderer.render(rp)                                # previous call left the plane: skipped
        dep[5, 5] = 2.0                                    # (the caller breaks its word, to make the skip observable)
        renderer.render(rp)
        assert dep[5, 5] == 2.0 and np.isposinf(dep).sum() == W * H - 1
        dep2 = np.full((H, W), -3.0, np.float32)           # another image: written
        rp2 = srb.RenderPass(srb.Image(col, W, H), srb.Image(dep2, W, H), v, i, srb.PrimitiveType.triangle, rp.transform, mesh_version=7)
        renderer.render(rp2)
        assert np.isposinf(dep2).all()
        renderer.depth_test = True                         # a depth-tested pass writes real depths ...
        renderer.render(rp2)
        assert np.isfinite(dep2).any()
        renderer.depth_test = False                        # ... so the next untested pass must write the plane again
        renderer.render(rp2)
        assert np.isposinf(dep2).all()
    finally:
        renderer.depth_untouched = False
        renderer.depth_test = False


def test_image_sizes_the_binning_fields_cannot_hold_are_refused(renderer):
    renderer.set_mesh(*S.cube())
    m = np.eye(4, dtype=np.float32)
    with pytest.raises(srb.SrError) as e:
        renderer.render_views(m[None], 600_000, 1, want_depth=False)
    assert e.value.code == srb.renderer.SR_ERR_ARG
    renderer.render_views(m[None], 300_000, 2, want_depth=False)          # 18,750 tile columns: fine


def test_resolve_kernel_with_unaligned_views_and_pointers():
    """Advisor finding r01: the 4-pixels-per-thread resolve needs 16-byte aligned planes; a batch of odd-sized views puts
    every second view on a 4-byte boundary, and a caller may pass a slice."""
    import torch
    W = H = 2049                                             # odd pixel count >= 4 Mi: the wide resolve is eligible
    v, i = S.cube_sphere(500)                                # 3M triangles of ~2 pixels: no binning (k_resolve path)
    views = S.fibonacci_views(4)[:2]
    with srb.CudaRenderer(0) as r:
        r.set_mesh(v, i)
        dev = torch.device("cuda", 0)
        color = torch.zeros((2 * H * W + 1, 4), dtype=torch.uint8, device=dev)
        depth = torch.zeros(2 * H * W + 1, dtype=torch.float32, device=dev)
        for off in (0, 1):                                   # off = 1: base pointers only 4-byte aligned
            for _ in range(2):                               # second pass runs the speculative no-binning frame
                r.render_device(views, W, H, color[off:].data_ptr(), depth[off:].data_ptr())
            assert r.stats()["binned_tris"] == 0
            got_c = color[off:off + 2 * H * W].cpu().numpy().reshape(2, H, W, 4)
            got_d = depth[off:off + 2 * H * W].cpu().numpy().reshape(2, H, W)
            for k in range(2):
                c, d = oracle_images(v, i, views[k], W, H)
                assert np.array_equal(got_c[k], c) and np.array_equal(got_d[k].view(np.uint32), d.view(np.uint32))


def test_frame_epoch_wrap_in_batched_frames():
    """tile_done stamps are compared with a 32-bit frame epoch: start three frames before the wrap."""
    code = r"""
import numpy as np, software_renderer_b200 as srb
from oracle import oracle as O
from software_renderer_b200 import scene as S
W, H = 480, 270
v, i = S.triangle_soup(70_000, W, H, seed=4)
with srb.CudaRenderer(0) as r:
    r.set_mesh(v, i)
    batched = 0
    for k in range(8):
        m = S.soup_transform(k)
        c, d = r.render_views(m[None], W, H)
        batched += r.stats()["batches_run"] > 0
        rc, oc, od, _ = O.render(v, i, m, W, H)
        assert np.array_equal(c[0], oc) and np.array_equal(d[0].view(np.uint32), od.view(np.uint32)), k
    assert batched >= 6
print("wrap ok")
"""
    env = dict(os.environ, SR_EPOCH0="0xFFFFFFFD", SR_BATCHES="1", PYTHONPATH=ROOT)
    r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, cwd=ROOT)
    assert r.returncode == 0 and "wrap ok" in r.stdout, r.stdout + r.stderr


def test_gather_without_a_communicator_is_an_error(renderer):
    sizes = (ctypes.c_size_t * 1)(16)
    rc = renderer._lib.sr_gather(renderer._ctx, None, None, sizes, 0, 0, None)
    assert rc == srb.renderer.SR_ERR_COMM


def test_host_fill_depth_any_alignment_and_size(renderer):
    """sr_host_fill_depth = clear(image: depthBuffer, with: .infinity) (Renderer.swift:206) on caller memory: the
    streaming-store loop has an unaligned head, 64-byte thread slices and a tail."""
    base = np.full(300_007, -2.0, np.float32)
    for off, n in ((0, 300_007), (1, 1), (3, 17), (5, 100_003), (2, 65_536), (7, 0)):
        base[:] = -2.0
        renderer.host_fill_depth(base[off:off + n])
        assert np.isposinf(base[off:off + n]).all()
        assert (base[:off] == -2.0).all() and (base[off + n:] == -2.0).all()       # nothing outside the range is touched
