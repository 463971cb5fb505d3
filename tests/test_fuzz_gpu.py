"""Randomised GPU parity sweep: random image sizes, triangle size mixes, shared indices, perspective transforms and
both visibility modes, repeated under every host policy the library can be forced into (slab / compact tile lists,
tiny slab segments, batched / single frames, with and without the no-binning speculation).  Every frame must match the
CPU oracle bit for bit; each renderer context keeps its speculative state across the sweep, so the policies are also
exercised in the order "guess from the previous frame, verify, re-run"."""
import os

import numpy as np
import pytest

import software_renderer_b200 as srb
from oracle import oracle as O
from software_renderer_b200 import scene as S

pytestmark = pytest.mark.gpu

POLICIES = [
    {},
    {"SR_SLAB": "0"},
    {"SR_SLAB_CAP": "2"},
    {"SR_BATCHES": "3"},
    {"SR_BATCHES": "3", "SR_SLAB": "0"},
    {"SR_BATCHES": "3", "SR_LBOX": "0"},        # batched frames without the box pre-pass (what frames wider than 4080 pixels get)
    {"SR_NOBIN": "0"},
    {"SR_ZPASS": "1"},                                             # experiment knobs that stay in the tree: two-pass depth frames,
    {"SR_BATCHES": "3", "SR_LBOX": "0", "SR_COMPACT": "1"},        # ... the compacting late k_setup without box words
    {"SR_PDL": "0"},
]


def scene(rng, dense=False):
    if rng.random() < 0.2:                      # a few big frames: many tiles, several batches, tile retirement
        W, H = int(rng.integers(1500, 2600)), int(rng.integers(900, 1700))
    else:
        W, H = int(rng.integers(1, 700)), int(rng.integers(1, 500))
    n = int(rng.integers(1, 400 if W > 1000 else 1500))       # the oracle pays per covered pixel
    kind = rng.integers(0, 4)
    if dense:            # enough triangles for index-descending batches (>= 2^16): mostly small, 2 % large on top
        n, kind = 70000, 4
    centre = rng.uniform(-1.2, 1.2, size=(n, 1, 3))
    if kind == 4:
        ext = np.where(rng.random(size=(n, 1, 1)) < 0.02, rng.uniform(0.2, 1.5, size=(n, 1, 1)), rng.uniform(0.0, 12.0 / max(W, H), size=(n, 1, 1)))
    elif kind == 0:        # tiny: a pixel or two
        ext = rng.uniform(0.0, 3.0 / max(W, H), size=(n, 1, 1))
    elif kind == 1:      # small: up to a tile
        ext = rng.uniform(0.0, 20.0 / max(W, H), size=(n, 1, 1))
    elif kind == 2:      # mixed, log-uniform from sub-pixel to screen-filling
        ext = np.exp(rng.uniform(np.log(1e-3), np.log(2.0), size=(n, 1, 1)))
    else:                # large (fewer of them: the oracle pays per covered pixel)
        n = min(n, 300)
        centre = centre[:n]
        ext = rng.uniform(0.2, 2.5, size=(n, 1, 1))
    xyz = (centre + rng.uniform(-1, 1, size=(n, 3, 3)) * ext).reshape(-1, 3)
    xyz[:, 2] = rng.uniform(-0.4, 1.0, size=3 * n)
    if rng.random() < 0.3:                      # snap to a coarse grid: equal y rows, coincident corners, depth ties
        xyz = np.round(xyz * 16) / 16
    rgb = rng.uniform(-0.3, 1.3, size=(3 * n, 3))
    v = S.pack_vertices(xyz.astype(np.float32), rgb.astype(np.float32))
    if rng.random() < 0.5:                      # shared / repeated vertices
        i = rng.integers(0, 3 * n, size=3 * n).astype(np.int64)
    else:
        i = np.arange(3 * n, dtype=np.int64)
    m = np.eye(4, dtype=np.float32)
    if rng.random() < 0.5:                      # App.swift:176-181 style perspective row, w stays positive
        m[3, 2] = rng.uniform(0.0, 0.6)
        m[3, 3] = 1.0
        m[0, 0] = rng.uniform(0.5, 1.5)
        m[1, 1] = rng.uniform(0.5, 1.5)
    return v, i, m, W, H


@pytest.fixture(scope="module")
def cases():
    """Scenes and their oracle images, computed once and shared by every policy."""
    rng = np.random.default_rng(0xF022)
    out = []
    for k in range(48):
        v, i, m, W, H = scene(rng, dense=k in (7, 21, 33))
        want = {}
        for dt in (False, True):
            rc, rc_c, rc_d, _ = O.render(v, i, m, W, H, flags=O.DEPTH_LESS if dt else 0)
            assert rc == O.OK
            want[dt] = (rc_c, rc_d)
        out.append((v, i, m, W, H, want))
    return out


@pytest.mark.parametrize("policy", POLICIES, ids=lambda p: ",".join(f"{k}={v}" for k, v in p.items()) or "default")
def test_fuzz_all_policies(policy, cases):
    # SR_POISON=1: every fresh device allocation is filled with 0xFF, so a read of a list entry, record or counter
    # that no kernel wrote faults (or shows) instead of passing by luck; allocations happen during the sweep
    env = dict(policy, SR_POISON="1")
    os.environ.update(env)
    try:
        r = srb.CudaRenderer(0)
        for case, (v, i, m, W, H, want) in enumerate(cases):
            for dt in (False, True):
                r.depth_test = dt
                rp = srb.RenderPass(srb.Image.color(W, H), srb.Image.depth(W, H), v, i, srb.PrimitiveType.triangle, m)
                rp.color_buffer.array[:] = 0xA5
                rp.depth_buffer.array[:] = 3.0
                r.render(rp)
                rc_c, rc_d = want[dt]
                bad = (rp.color_buffer.array != rc_c).any(-1)
                assert not bad.any(), f"case {case} {W}x{H} depth={dt}: {int(bad.sum())} colour pixels differ, first {np.argwhere(bad)[0]}"
                assert np.array_equal(rp.depth_buffer.array.view(np.uint32), rc_d.view(np.uint32)), f"case {case} depth={dt}: depth differs"
        r.depth_test = False
        r.close()
    finally:
        for k in env:
            del os.environ[k]


def _view(rng, scale):
    """rotation * scale, pushed back along z, then the w = z + 1 projection of App.swift:176-181."""
    q = rng.normal(size=4)
    q /= np.linalg.norm(q)
    w, x, y, z = q
    R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                  [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                  [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])
    m = np.eye(4)
    m[:3, :3] = R * scale
    m[2, 3] = 1.5
    p = np.eye(4)
    p[3, 2] = 1.0
    return (p @ m).astype(np.float32)


@pytest.fixture(scope="module")
def view_cases():
    """One shared-vertex mesh seen from 14 cameras whose zoom spans sub-pixel triangles to a few screen-filling ones
    (some corners behind the eye: negative w, still finite).  Consecutive views want different host policies."""
    rng = np.random.default_rng(0xCA3E)
    v, i = S.cube_sphere(40)
    W, H = 333, 251
    views = np.stack([_view(rng, s) for s in np.exp(rng.uniform(np.log(0.1), np.log(8.0), size=14))])
    want = {}
    for dt in (False, True):
        outs = [O.render(v, i, m, W, H, flags=O.DEPTH_LESS if dt else 0) for m in views]
        want[dt] = outs
    return v, i, views, W, H, want


@pytest.mark.parametrize("policy", POLICIES, ids=lambda p: ",".join(f"{k}={v}" for k, v in p.items()) or "default")
def test_fuzz_views_all_policies(policy, view_cases):
    v, i, views, W, H, want = view_cases
    env = dict(policy, SR_POISON="1")
    os.environ.update(env)
    try:
        r = srb.CudaRenderer(0)
        r.set_mesh(v, i)
        for dt in (False, True):
            r.depth_test = dt
            for rounds in range(2):                      # second round starts from the policies the first one left
                color = np.full((len(views), H, W, 4), 0x3C, np.uint8)
                depth = np.full((len(views), H, W), -1.0, np.float32)
                skipped = sum(o[3]["domain_skipped"] for o in want[dt])
                try:
                    r.render_views(views, W, H, color_out=color, depth_out=depth)
                    assert skipped == 0
                except srb.DomainError:
                    assert skipped > 0
                for k, (rc, rc_c, rc_d, _) in enumerate(want[dt]):
                    bad = (color[k] != rc_c).any(-1)
                    assert not bad.any(), f"view {k} depth={dt} round {rounds}: {int(bad.sum())} colour pixels differ"
                    assert np.array_equal(depth[k].view(np.uint32), rc_d.view(np.uint32)), f"view {k} depth={dt}: depth differs"
        r.depth_test = False
        r.close()
    finally:
        for k in env:
            del os.environ[k]
