"""sr_points_no_alloc: the allocation-free entry point (no counterpart in the reference).
Called through the C ABI and compared with the EXISTING entry point for the same pass
(sr_render_device, PrimitiveType.vertices) - bit-exact, the bar of every other test here - on two calls in a row (the
static scratch array is shared between calls and must come back clean).  This file is collected last on purpose."""
import ctypes

import numpy as np
import pytest

import software_renderer_b200 as srb
from software_renderer_b200 import scene as S

pytestmark = pytest.mark.gpu
MAX_PIXELS = 1 << 20          # SR_NOALLOC_MAX_PIXELS (include/sr_cuda.h)


def colmajor(m):
    return np.ascontiguousarray(np.asarray(m, np.float32).T).reshape(16)


def no_alloc(lib, v_dev, nverts, i_dev, nidx, m, W, H, color, depth, stream=None):
    mm = colmajor(m)
    return lib.sr_points_no_alloc(0, ctypes.c_void_p(v_dev.data_ptr()), nverts, ctypes.c_void_p(i_dev.data_ptr()), nidx,
                                  mm.ctypes.data, W, H, ctypes.c_void_p(color.data_ptr()),
                                  ctypes.c_void_p(depth.data_ptr()) if depth is not None else None,
                                  ctypes.c_void_p(stream) if stream else None)


def test_no_alloc_points_match_the_context_path_twice_in_a_row():
    import torch
    lib = srb.load_library()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    W, H = 640, 360
    v, i = S.icosphere(4)                                   # 2,562 shared vertices, 15,360 corners: many collide on a pixel
    i = np.ascontiguousarray(i, np.int64)
    v_dev, i_dev = torch.from_numpy(v).to(dev), torch.from_numpy(i).to(dev)
    with srb.CudaRenderer(0) as r:
        r.set_mesh(v, i)
        for k, t in enumerate((0.3, 1.1)):                  # two calls in a row, two different transforms
            m = S.app_transform(t)
            want_c = torch.full((H, W, 4), 0x5A, dtype=torch.uint8, device=dev)
            want_d = torch.full((H, W), -7.0, dtype=torch.float32, device=dev)
            r.render_device(m, W, H, want_c.data_ptr(), want_d.data_ptr(), primitive_type=srb.PrimitiveType.vertices)
            got_c = torch.full((H, W, 4), 0xA5, dtype=torch.uint8, device=dev)       # stale contents must be overwritten
            got_d = torch.full((H, W), 3.0, dtype=torch.float32, device=dev)
            free_before = torch.cuda.mem_get_info(0)[0]
            rc = no_alloc(lib, v_dev, v.shape[0], i_dev, i.size, m, W, H, got_c, got_d)
            assert rc == 0, lib.sr_points_no_alloc_last_error()
            torch.cuda.synchronize(dev)
            if k == 1:                                      # the first call may load the module lazily; the second allocates nothing
                assert torch.cuda.mem_get_info(0)[0] == free_before
            assert (want_c[..., 3] != 0).sum().item() > 500                       # the pass drew something
            assert torch.equal(got_c, want_c), f"call {k}: {int((got_c != want_c).any(-1).sum())} pixels differ"
            assert torch.equal(got_d.view(torch.int32), want_d.view(torch.int32))
        # colour only (depth NULL), on a caller's stream
        st = torch.cuda.Stream(device=dev)
        got_c = torch.zeros((H, W, 4), dtype=torch.uint8, device=dev)
        assert no_alloc(lib, v_dev, v.shape[0], i_dev, i.size, m, W, H, got_c, None, st.cuda_stream) == 0
        st.synchronize()
        assert torch.equal(got_c, want_c)


def test_no_alloc_limits_and_edge_cases():
    import torch
    lib = srb.load_library()
    dev = torch.device("cuda", 0)
    v, i = S.cube()
    i = np.ascontiguousarray(i, np.int64)
    v_dev, i_dev = torch.from_numpy(v).to(dev), torch.from_numpy(i).to(dev)
    m = S.app_transform(0.2)
    big = torch.zeros(8, dtype=torch.uint8, device=dev)
    # one pixel beyond the static scratch: refused with a clear message, nothing launched
    rc = no_alloc(lib, v_dev, v.shape[0], i_dev, i.size, m, 1025, 1024, big, None)
    assert rc == srb.renderer.SR_ERR_ARG
    msg = lib.sr_points_no_alloc_last_error().decode()
    assert "exceed" in msg and str(MAX_PIXELS) in msg
    assert no_alloc(lib, v_dev, v.shape[0], i_dev, i.size, m, 0, 16, big, None) == srb.renderer.SR_ERR_ARG
    # exactly the maximum is accepted; no corners at all = the clears only (R:205-206)
    W = H = 1024
    c = torch.full((H, W, 4), 9, dtype=torch.uint8, device=dev)
    d = torch.zeros((H, W), dtype=torch.float32, device=dev)
    assert no_alloc(lib, v_dev, v.shape[0], i_dev, 0, m, W, H, c, d) == 0
    torch.cuda.synchronize(dev)
    assert int(c.max()) == 0 and bool(torch.isposinf(d).all())
    # an out-of-range index plots nothing and corrupts nothing; a trailing partial primitive is ignored (R:221)
    bad = i.copy()
    bad[0], bad[1] = 10 ** 9, -5
    bad_dev = torch.from_numpy(np.concatenate([bad, [0, 1]]).astype(np.int64)).to(dev)
    assert no_alloc(lib, v_dev, v.shape[0], bad_dev, bad.size + 2, m, 320, 240, c, d) == 0
    torch.cuda.synchronize(dev)
    # ... and the shared array came back clean: the same pass through the context path equals a fresh no-alloc call
    with srb.CudaRenderer(0) as r:
        r.set_mesh(v, i)
        want = torch.zeros((240, 320, 4), dtype=torch.uint8, device=dev)
        r.render_device(m, 320, 240, want.data_ptr(), None, primitive_type=srb.PrimitiveType.vertices)
    got = torch.zeros((240, 320, 4), dtype=torch.uint8, device=dev)
    assert no_alloc(lib, v_dev, v.shape[0], i_dev, i.size, m, 320, 240, got, None) == 0
    torch.cuda.synchronize(dev)
    assert torch.equal(got, want)
