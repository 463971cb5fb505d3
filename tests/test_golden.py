"""Golden vectors (tests/golden/golden.npz, made by tests/golden/make_golden.py from two independent restatements of
Renderer.swift that had to agree bit for bit).  CPU: both forms of the C++ oracle and the pure-Python restatement
must still reproduce them.  GPU: the CUDA path, through the C ABI, must reproduce them too."""
import os

import numpy as np
import pytest

from oracle import literal_py as P
from oracle import oracle as O

GOLDEN = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden.npz"))
NAMES = [str(n) for n in GOLDEN["names"]]


def case(name):
    W, H, prim, depth, skipped = (int(x) for x in GOLDEN[f"{name}/meta"])
    return (GOLDEN[f"{name}/vertices"], GOLDEN[f"{name}/indices"], GOLDEN[f"{name}/transform"], W, H, prim, bool(depth),
            skipped, GOLDEN[f"{name}/color"], GOLDEN[f"{name}/depth"])


def same(c, d, gc, gd):
    bad = (c != gc).any(-1)
    assert not bad.any(), f"{int(bad.sum())} colour pixels differ, first at {np.argwhere(bad)[0]}"
    assert np.array_equal(d.view(np.uint32), gd.view(np.uint32)), "depth differs"


def test_fixture_is_complete():
    assert len(NAMES) >= 17
    assert {"kat1_64", "cube_t0", "soup60_depth", "tie_depth", "domain_skip", "ico1_vertices", "ico1_line"} <= set(NAMES)


@pytest.mark.parametrize("literal", [False, True])
@pytest.mark.parametrize("name", NAMES)
def test_oracle_reproduces_golden(name, literal):
    v, i, m, W, H, prim, depth, skipped, gc, gd = case(name)
    flags = (O.LITERAL if literal else 0) | (O.DEPTH_LESS if depth else 0)
    rc, c, d, st = O.render(v, i, m, W, H, primitive_type=prim, flags=flags)
    assert rc == (O.ERR_DOMAIN if skipped else O.OK)
    assert st["domain_skipped"] == skipped
    same(c, d, gc, gd)


@pytest.mark.parametrize("name", ["kat2_64", "cube_t13_depth", "tiny40", "tie_depth", "domain_skip"])
def test_python_restatement_reproduces_golden(name):
    v, i, m, W, H, prim, depth, skipped, gc, gd = case(name)
    c, d, sk = P.render(v, i, m, W, H, primitive_type=prim, depth_less=depth)
    assert sk == skipped
    same(c, d, gc, gd)


def test_kat_values_inside_golden():
    """KAT-1 at 64x64: rows 0..31 span [32-y, 32+y], row 32 collapses to the single pixel (0, 32)."""
    *_, gc, gd = case("kat1_64")
    cov = gc[..., 3] > 0
    assert cov.sum() == 32 * 32 + 1
    assert np.nonzero(cov[32])[0].tolist() == [0]
    assert gc[0, 32].tolist() == [0, 0, 255, 255] and gc[32, 0].tolist() == [255, 0, 0, 255]
    assert np.isinf(gd).all()


@pytest.mark.gpu
@pytest.mark.parametrize("name", NAMES)
def test_cuda_reproduces_golden(renderer, name):
    import software_renderer_b200 as srb
    v, i, m, W, H, prim, depth, skipped, gc, gd = case(name)
    renderer.depth_test = depth
    rp = srb.RenderPass(srb.Image.color(W, H), srb.Image.depth(W, H), v, i, srb.PrimitiveType(prim), m)
    try:
        if skipped:
            with pytest.raises(srb.DomainError):
                renderer.render(rp)
        else:
            renderer.render(rp)
    finally:
        renderer.depth_test = False
    same(rp.color_buffer.array, rp.depth_buffer.array, gc, gd)
