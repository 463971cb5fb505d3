"""The SR_SIMD_FMA variant (Apple's simd_muladd fused, as arm64 compiles Renderer.swift:160 and :91) of the oracle and
of the CUDA library.  CPU: both oracle builds run and differ only where rounding may move a result; GPU: the fused CUDA
build is bit-exact against the fused oracle, like the default pair."""
import numpy as np
import pytest

import software_renderer_b200 as srb
from oracle import oracle as O
from software_renderer_b200 import scene as S


def both(v, i, m, W, H, depth):
    out = []
    for fma in (False, True):
        rc, c, d, _ = O.render(v, i, m, W, H, flags=O.DEPTH_LESS if depth else 0, simd_fma=fma)
        assert rc == O.OK
        out.append((c, d))
    return out


def test_matvec_variants_differ_in_the_last_bit_only():
    rng = np.random.default_rng(5)
    m = S.app_transform(0.37)
    worst = 0.0
    differs = 0
    for _ in range(2000):
        p = rng.uniform(-1, 1, 3).astype(np.float32)
        a = np.ascontiguousarray(O.transform_vertex(m, p, 1920, 1080))
        out = np.empty(3, np.float32)
        mm = np.ascontiguousarray(np.asarray(m, np.float32).T)
        O.lib(True).sr_oracle_transform_vertex(mm.ctypes.data, p.ctypes.data, 1920, 1080, out.ctypes.data)
        differs += int((a.view(np.uint32) != out.view(np.uint32)).any())
        worst = max(worst, float((np.abs(a - out) / np.maximum(1.0, np.abs(a))).max()))
    assert differs > 0, "the two builds are the same arithmetic: SR_SIMD_FMA is not in effect"
    assert worst < 1e-4            # screen coordinates move by rounding noise (relative), not by pixels


@pytest.mark.parametrize("depth", [False, True])
def test_oracle_variants_on_the_cube_scene(depth):
    v, i = S.cube()
    (c0, d0), (c1, d1) = both(v, i, S.app_transform(7 / 60.0), 800, 600, depth)
    assert np.array_equal(c0[..., 3] != 0, c1[..., 3] != 0) or ((c0[..., 3] != 0) != (c1[..., 3] != 0)).sum() < 50
    colour = (c0 != c1).any(-1).sum()
    assert colour < 0.01 * 800 * 600        # measured: ~1,400 of 480,000 pixels, each by one LSB


def test_oracle_variants_on_an_icosphere():
    v, i = S.icosphere(4)
    (c0, d0), (c1, d1) = both(v, i, S.app_transform(0.5), 640, 360, True)
    assert (c0 != c1).any(-1).sum() < 0.01 * 640 * 360
    assert np.isfinite(d1[c1[..., 3] != 0]).all()


def test_fused_cuda_library_exports_the_abi():
    lib = srb.load_library(simd_fma=True)
    for name in srb.ABI_SYMBOLS:
        getattr(lib, name)


@pytest.fixture(scope="module")
def fused():
    r = srb.CudaRenderer(0, simd_fma=True)
    yield r
    r.close()


@pytest.mark.gpu
@pytest.mark.parametrize("depth", [False, True])
def test_fused_cuda_equals_fused_oracle(fused, depth):
    rng = np.random.default_rng(11)
    cases = [(S.icosphere(5), S.app_transform(0.4), 1280, 720),
             (S.cube(), S.app_transform(1.3), 800, 600),
             (S.triangle_soup(60_000, 960, 540, seed=3), S.soup_transform(2), 960, 540),
             (S.cube_sphere(120), S.fibonacci_views(8)[5], 512, 512)]
    fused.depth_test = depth
    for (v, i), m, W, H in cases:
        rp = srb.RenderPass(srb.Image.color(W, H), srb.Image.depth(W, H), v, i, srb.PrimitiveType.triangle, m)
        for _ in range(2):                      # second frame: the adaptive launch policies have settled
            rp.color_buffer.array[:] = rng.integers(0, 255)
            fused.render(rp)
        rc, c, d, _ = O.render(v, i, m, W, H, flags=O.DEPTH_LESS if depth else 0, simd_fma=True)
        assert rc == O.OK
        assert np.array_equal(rp.color_buffer.array, c)
        assert np.array_equal(rp.depth_buffer.array.view(np.uint32), d.view(np.uint32))
    fused.depth_test = False


@pytest.mark.gpu
def test_fused_and_default_cuda_differ_like_the_oracles(fused, renderer):
    """The two CUDA builds really are different arithmetic (the flag reaches the kernels)."""
    v, i = S.cube()
    m = S.app_transform(7 / 60.0)
    out = []
    for r in (renderer, fused):
        rp = srb.RenderPass(srb.Image.color(800, 600), srb.Image.depth(800, 600), v, i, srb.PrimitiveType.triangle, m)
        r.render(rp)
        out.append(rp.color_buffer.array.copy())
    assert (out[0] != out[1]).any()
