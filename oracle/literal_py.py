"""Second, independent CPU restatement of Renderer.render(renderPass:) in pure Python (numpy float32 scalars).

TEST INFRASTRUCTURE ONLY.  It exists to pin oracle/sr_oracle.cpp: it was written from the Swift source alone
(/root/reference/renderer/Renderer.swift, cited per function as R:line) without looking at the C++ restatement,
keeps the reference's per-pixel structure (the 2x2 inverse is recomputed for every pixel; off-screen rows and pixels,
which the reference rejects one by one at R:246-250, are stepped over), and is only usable on small cases.  tests/golden/make_golden.py runs both
restatements on the same inputs and refuses to write fixtures unless they agree bit for bit.

Apple `simd` operations that are not in the reference tree are defined exactly as in oracle/sr_oracle.h:
  matrix * vector : column accumulation, every * and + rounded separately (no FMA)
  float2x2.inverse: adjugate / det, det = m00*m11 - m01*m10, four true divisions, no singular early-out
  simd_clamp      : min(max(v, lo), hi) with fmax/fmin NaN handling
"""
from __future__ import annotations

import math

import numpy as np

F = np.float32
TRIANGLE, LINE, VERTICES = 0, 1, 2
_LIMIT = F(2.0 ** 30)


class DomainTrap(Exception):
    """The reference would trap here (Int(Float) of NaN/inf/out-of-range, R:251, R:271, R:298)."""


def _fmax(a, b):          # C fmaxf: a NaN operand is ignored
    if math.isnan(a):
        return b
    if math.isnan(b):
        return a
    return a if a > b else b


def _fmin(a, b):
    if math.isnan(a):
        return b
    if math.isnan(b):
        return a
    return a if a < b else b


def _trunc(v) -> int:     # Int(Float) / simd_long2(float2): toward zero, traps outside the domain
    if not (abs(v) < _LIMIT):
        raise DomainTrap()
    return int(v)


def apply(m_cols, xyz):
    """Vertex.apply (R:159-163).  m_cols[c][r] like matrix_float4x4.columns."""
    v = (xyz[0], xyz[1], xyz[2], F(1))
    acc = [F(m_cols[0][r] * v[0]) for r in range(4)]
    for c in (1, 2, 3):
        acc = [F(acc[r] + F(m_cols[c][r] * v[c])) for r in range(4)]
    with np.errstate(all="ignore"):                             # w = 0 gives inf/NaN, caught by the domain rule
        return F(acc[0] / acc[3]), F(acc[1] / acc[3]), F(acc[2] / acc[3])


def converted_to_screen(xyz, width, height):
    """convertedToScreen (R:165-172): .rounded applies to the (W, H) vector only."""
    ux = F(F(xyz[0] * F(0.5)) + F(0.5))
    uy = F(F(xyz[1] * F(-0.5)) + F(0.5))
    return F(ux * F(width)), F(uy * F(height)), xyz[2]


def interpolate(values, t):
    """R:467-494, Swift Int arithmetic (Python ints; truncating division)."""
    base = 0
    if len(values) == 3:
        if t >= values[2][1]:
            base = 2
        elif t >= values[1][1]:
            base = 1
    nxt = base + 1
    start = values[base]
    if nxt >= len(values):
        return start[0]
    end = values[nxt]
    diff = end[0] - start[0]
    dy = end[1] - start[1]
    if dy == 0:
        return start[0]
    nt = t - start[1]
    num = diff * nt
    q = abs(num) // abs(dy)
    if (num < 0) != (dy < 0):
        q = -q
    return start[0] + q


def ws(a, b, c, x, y):
    """Triangle.ws(xy:) with T() recomputed, as setPixel does per pixel (R:88-100, R:251-252)."""
    afx, afy = F(F(a[0]) + F(0.5)), F(F(a[1]) + F(0.5))
    bfx, bfy = F(F(b[0]) + F(0.5)), F(F(b[1]) + F(0.5))
    cfx, cfy = F(F(c[0]) + F(0.5)), F(F(c[1]) + F(0.5))
    m00, m10 = F(afx - cfx), F(afy - cfy)        # column 0
    m01, m11 = F(bfx - cfx), F(bfy - cfy)        # column 1
    with np.errstate(all="ignore"):
        det = F(F(m00 * m11) - F(m01 * m10))
        i00, i10 = F(m11 / det), F(F(-m10) / det)   # inverse column 0
        i01, i11 = F(F(-m01) / det), F(m00 / det)   # inverse column 1
        px, py = F(F(x) + F(0.5)), F(F(y) + F(0.5))
        dx, dy = F(px - cfx), F(py - cfy)
        wa = F(F(i00 * dx) + F(i01 * dy))
        wb = F(F(i10 * dx) + F(i11 * dy))
        wc = F(F(F(1) - wa) - wb)
    return wa, wb, wc


def pack(rgb):
    """Pixel(float3:) -> bytes b, g, r, a (R:117-128)."""
    out = []
    for v in (rgb[2], rgb[1], rgb[0], F(1)):
        cl = _fmin(_fmax(float(v), 0.0), 1.0)
        out.append(int(F(F(cl) * F(255))))
    return out


def render(vertices, indices, transform, width, height, primitive_type=TRIANGLE, depth_less=False):
    """Returns (color[H,W,4] uint8, depth[H,W] float32, domain_skipped).  `transform` is (4,4), M @ v convention.
    A primitive the reference would trap on is skipped and counted (the documented error convention)."""
    verts = np.asarray(vertices, dtype=np.float32)
    idx = [int(k) for k in np.asarray(indices).reshape(-1)]
    m = np.asarray(transform, dtype=np.float32)
    m_cols = [[F(m[r, c]) for r in range(4)] for c in range(4)]
    color = np.zeros((height, width, 4), np.uint8)               # R:205
    depth = np.full((height, width), np.inf, np.float32)        # R:206
    per = 2 if primitive_type == LINE else 3                     # R:179-188
    skipped = 0
    for p in range(len(idx) // per):                             # R:221-229
        corner = [verts[idx[p * per + k]] for k in range(per)]
        tv = [(apply(m_cols, (F(v[0]), F(v[1]), F(v[2]))), (F(v[4]), F(v[5]), F(v[6]))) for v in corner]
        if primitive_type == LINE:                               # R:289-293 draws nothing
            continue
        if primitive_type == VERTICES:                           # R:295-302; a corner the reference would trap on
            for xyz, col in tv:                                  # is skipped and counted on its own
                s = converted_to_screen(xyz, width, height)
                try:
                    x, y = _trunc(s[0]), _trunc(s[1])
                except DomainTrap:
                    skipped += 1
                    continue
                if 0 <= x < width and 0 <= y < height:           # R:30-36
                    color[y, x] = pack(col)
            continue
        try:
            scr = [converted_to_screen(xyz, width, height) for xyz, _ in tv]      # R:241-243
            order = sorted(range(3), key=lambda k: scr[k][1])                     # stable, strict < on float y (R:271)
            ints = [(_trunc(s[0]), _trunc(s[1])) for s in scr]
        except DomainTrap:
            skipped += 1
            continue
        s0, s1, s2 = (ints[k] for k in order)
        left, right = [s0, s1, s2], [s0, s2]                     # R:272-273
        za, zb, zc = scr[0][2], scr[1][2], scr[2][2]
        for y in range(s0[1], s2[1] + 1):                        # R:275
            lx, rx = interpolate(left, y), interpolate(right, y)
            if lx > rx:
                lx, rx = rx, lx
            if not (0 <= y < height):
                continue                                         # every pixel of the row fails R:246-250
            for x in range(max(lx, 0), min(rx, width - 1) + 1):  # R:281; off-screen x fail R:246-250 one by one
                w = ws(ints[0], ints[1], ints[2], x, y)          # original a, b, c order (R:251)
                with np.errstate(all="ignore"):
                    if depth_less:                               # R:257-261, uncommented
                        d = F(F(F(za * w[0]) + F(zb * w[1])) + F(zc * w[2]))
                        if not (d < depth[y, x]):
                            continue
                        depth[y, x] = d
                    rgb = [F(F(F(tv[0][1][k] * w[0]) + F(tv[1][1][k] * w[1])) + F(tv[2][1][k] * w[2])) for k in range(3)]
                color[y, x] = pack(rgb)                          # R:263-268
    return color, depth, skipped
