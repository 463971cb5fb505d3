"""Legacy 2-D immediate-mode API (Renderer.swift:376-465).
CPU: the oracle against hand-derived answers and a pure-Python reading of the Swift source.
GPU: sr_draw2d (CUDA kernels through the C ABI) bit for bit against the oracle."""
import math

import numpy as np
import pytest

from oracle import oracle as O
from software_renderer_b200.renderer import Cmd2D, pack_cmds

F = np.float32


def run_oracle(cmds, W, H, fill=0x11):
    img = np.full((H, W, 4), fill, np.uint8)
    rc = O.draw2d(pack_cmds(cmds), len(cmds), img)
    return rc, img


def py_reference(cmds, W, H, fill=0x11):
    """Straight reading of R:376-465 with numpy float32 scalars (independent of the C++ restatement)."""
    from oracle import literal_py as P
    img = np.full((H, W, 4), fill, np.uint8)
    sg = lambda v: (v > 0) - (v < 0)
    for c in cmds:
        p, col = list(c.p), list(c.color)
        if c.kind == Cmd2D.RECT:
            x0, y0, w, h = p[:4]
            for y in range(y0, y0 + w, sg(w)):
                for x in range(x0, x0 + h, sg(h)):
                    img[y, x] = col
        elif c.kind == Cmd2D.CIRCLE:
            cx, cy, r = p[:3]
            for dy in range(-r, r, sg(r)):
                for dx in range(-r, r, sg(r)):
                    px, py = F(F(dx) + F(0.5)), F(F(dy) + F(0.5))
                    l = F(np.sqrt(F(F(px * px) + F(py * py))))
                    rf = F(r)
                    if l < F(rf - F(2)):
                        d = F(1)
                    else:
                        e0, e1 = F(rf + F(2)), F(rf - F(2))
                        t = F(min(max(F(F(l - e0) / F(e1 - e0)), F(0)), F(1)))
                        d = F(min(max(F(F(t * t) * F(F(3) - F(F(2) * t))), F(0)), F(1)))
                    img[cy + dy, cx + dx] = [int(F(d * F(255))), 0, 0, 255]
        elif c.kind == Cmd2D.LINE:
            x0, y0, x1, y1 = p[:4]
            dx, dy = x1 - x0, y1 - y0
            steps = max(abs(dx), abs(dy))
            if steps:
                xs, ys = F(F(dx) / F(steps)), F(F(dy) / F(steps))
                x, y = F(x0), F(y0)
                for _ in range(steps):
                    rx = int(math.floor(abs(float(x)) + 0.5)) * (1 if x >= 0 else -1)     # .rounded(): ties away from zero
                    ry = int(math.floor(abs(float(y)) + 0.5)) * (1 if y >= 0 else -1)
                    if 0 <= rx < W and 0 <= ry < H:
                        img[ry, rx] = col
                    x, y = F(x + xs), F(y + ys)
        else:
            a, b, cc = (p[0], p[1]), (p[2], p[3]), (p[4], p[5])
            s = sorted([a, b, cc], key=lambda q: q[1])
            for y in range(s[0][1], s[2][1] + 1):
                l, r = P.interpolate(s, y), P.interpolate([s[0], s[2]], y)
                if l > r:
                    l, r = r, l
                for x in range(l, r + 1):
                    if 0 <= x < W and 0 <= y < H:
                        with np.errstate(all="ignore"):
                            w = P.ws(a, b, cc, x, y)
                            rgb = [F(F(F(F(w[0] * F(u0)) + F(w[1] * F(u1))) + F(w[2] * F(u2))) * F(1))
                                   for u0, u1, u2 in ((1, 0, 0), (0, 1, 0), (0, 0, 1))]
                        img[y, x] = P.pack(rgb)
    return img


def scene_cmds(seed, W, H, n=24):
    rng = np.random.default_rng(seed)
    cmds = []
    for k in range(n):
        kind = k % 4
        col = [int(v) for v in rng.integers(0, 256, size=4)]
        if kind == Cmd2D.RECT:
            w, h = int(rng.integers(1, H // 3)) * int(rng.choice([-1, 1])), int(rng.integers(1, W // 3)) * int(rng.choice([-1, 1]))
            y = int(rng.integers(max(0, -w), min(H, H - w) if w > 0 else H))
            x = int(rng.integers(max(0, -h), min(W, W - h) if h > 0 else W))
            cmds.append(Cmd2D.rect(x, y, w, h, col))
        elif kind == Cmd2D.CIRCLE:
            r = int(rng.integers(1, min(W, H) // 4)) * int(rng.choice([-1, 1]))
            a = abs(r)
            cmds.append(Cmd2D.circle(int(rng.integers(a, W - a)), int(rng.integers(a, H - a)), r))
        elif kind == Cmd2D.LINE:
            p = rng.integers(-20, max(W, H) + 20, size=4)
            cmds.append(Cmd2D.line(p[0], p[1], p[2], p[3], col))
        else:
            p = rng.integers(-15, max(W, H) + 15, size=6)
            cmds.append(Cmd2D.triangle((p[0], p[1]), (p[2], p[3]), (p[4], p[5])))
    return cmds


def test_rect_swaps_w_and_h_as_the_reference_does():
    rc, img = run_oracle([Cmd2D.rect(2, 1, 3, 5, (9, 8, 7, 6))], 12, 10)     # w = 3 rows, h = 5 columns
    assert rc == O.OK
    cov = (img == np.array([9, 8, 7, 6], np.uint8)).all(-1)
    assert cov.sum() == 15 and cov[1:4, 2:7].all()
    rc, img = run_oracle([Cmd2D.rect(6, 4, -2, -3, (1, 2, 3, 4))], 12, 10)    # negative extents walk backwards
    cov = (img == np.array([1, 2, 3, 4], np.uint8)).all(-1)
    assert cov.sum() == 6 and cov[3:5, 4:7].all()


def test_circle_writes_blue_only_and_is_smooth():
    rc, img = run_oracle([Cmd2D.circle(16, 16, 10)], 32, 32)
    assert rc == O.OK
    box = img[6:26, 6:26]
    assert (box[..., 1] == 0).all() and (box[..., 2] == 0).all() and (box[..., 3] == 255).all()
    assert box[10, 10, 0] == 255                       # centre: |p| < r - 2
    assert box[0, 0, 0] == 0                           # corner of the box: |p| > r + 2
    assert 0 < img[16, 16 + 9, 0] < 255                # on the smooth-step edge
    assert (img[0:6] == 0x11).all()                    # untouched rows keep their contents


def test_line_dda_and_triangle_weights():
    rc, img = run_oracle([Cmd2D.line(0, 0, 8, 4, (5, 6, 7, 8))], 10, 10)
    pts = np.argwhere((img == np.array([5, 6, 7, 8], np.uint8)).all(-1))
    assert len(pts) == 8 and [0, 0] in pts.tolist() and [4, 8] not in pts.tolist()      # `0..<steps`: the end point is not drawn
    rc, img = run_oracle([Cmd2D.triangle((32, 0), (64, 32), (0, 32))], 64, 64)          # KAT-1 geometry at 64x64
    cov = img[..., 3] == 255
    assert cov.sum() == 32 * 32 + 1                    # rows 0..31: [32-y, 32+y]; row 32 collapses to the pixel (0, 32)
    assert img[0, 32].tolist() == [0, 0, 255, 255]     # weight 1 on corner a -> red


def test_domain_errors_draw_nothing():
    for bad in (Cmd2D.rect(8, 8, 5, 5, (1, 1, 1, 1)), Cmd2D.rect(1, 1, 0, 3, (1, 1, 1, 1)), Cmd2D.circle(3, 3, 5), Cmd2D.circle(5, 5, 0)):
        rc, img = run_oracle([Cmd2D.rect(0, 0, 2, 2, (3, 3, 3, 3)), bad], 10, 10)
        assert rc == O.ERR_DOMAIN and (img == 0x11).all()


@pytest.mark.parametrize("seed", range(4))
def test_oracle_equals_python_reading(seed):
    W, H = 57, 41
    cmds = scene_cmds(seed, W, H)
    rc, img = run_oracle(cmds, W, H)
    assert rc == O.OK
    assert np.array_equal(img, py_reference(cmds, W, H))


@pytest.mark.gpu
@pytest.mark.parametrize("seed,size", [(10, (57, 41)), (11, (640, 360)), (12, (1920, 1080)), (13, (33, 200))])
def test_cuda_draw2d_equals_oracle(renderer, seed, size):
    W, H = size
    cmds = scene_cmds(seed, W, H, n=40)
    rc, want = run_oracle(cmds, W, H)
    assert rc == O.OK
    got = np.full((H, W, 4), 0x11, np.uint8)
    renderer.draw2d(cmds, got)
    bad = (got != want).any(-1)
    assert not bad.any(), f"{int(bad.sum())} pixels differ, first at {np.argwhere(bad)[0]}"
    assert renderer.stats()["gpu_launches"] == len(cmds)


@pytest.mark.gpu
def test_cuda_draw2d_errors(renderer):
    import software_renderer_b200 as srb
    img = np.full((10, 10, 4), 0x11, np.uint8)
    for bad in (Cmd2D.rect(8, 8, 5, 5, (1, 1, 1, 1)), Cmd2D.circle(3, 3, 5), Cmd2D.circle(5, 5, 0)):
        with pytest.raises(srb.SrError):
            renderer.draw2d([Cmd2D.rect(0, 0, 2, 2, (3, 3, 3, 3)), bad], img)
        assert (img == 0x11).all()
