#!/usr/bin/env python
"""Generates tests/golden/golden.npz — small input/output vectors for Renderer.render(renderPass:).

The reference (Swift + Apple simd) cannot run here and ships no tests or golden vectors, so the fixtures are produced
by TWO independent restatements of /root/reference/renderer/Renderer.swift run on the same inputs:
  * oracle/sr_oracle.cpp   (C++, literal form)          — the oracle the GPU parity tests use
  * oracle/literal_py.py   (pure Python, float32 scalars) — written from the Swift source alone
The script refuses to write the file unless both agree bit for bit on colour, depth and the skipped count.
Frozen outputs then guard the oracle (and, through `-m gpu`, the CUDA path) against silent changes.

    python tests/golden/make_golden.py          # rewrites golden.npz
"""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import literal_py as P          # noqa: E402
from oracle import oracle as O              # noqa: E402
from software_renderer_b200 import scene as S   # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden.npz")
I4 = np.eye(4, dtype=np.float32)


def tri(xy, z=(0, 0, 0), rgb=None):
    xyz = np.array([[x, y, zz] for (x, y), zz in zip(xy, z)], np.float32)
    return S.pack_vertices(xyz, np.eye(3, dtype=np.float32) if rgb is None else np.asarray(rgb, np.float32))


def cases():
    """name -> (vertices, indices, transform, W, H, primitive, depth_less)"""
    rng = np.random.default_rng(0x601D)
    out = {}
    # KAT-1 / KAT-2 geometry at reduced size
    out["kat1_64"] = (tri([(0, 1), (1, 0), (-1, 0)]), np.arange(3), I4, 64, 64, 0, False)
    out["kat2_64"] = (tri([(0, 0.5), (-0.5, -0.5), (0.5, -0.5)]), np.arange(3), I4, 64, 64, 0, False)
    # config 1: the rotating cube through the app's transform recipe (App.swift:169-183)
    v, i = S.cube()
    for k in (0, 13):
        out[f"cube_t{k}"] = (v, i, S.app_transform(k / 60.0), 80, 60, 0, False)
    out["cube_t13_depth"] = (v, i, S.app_transform(13 / 60.0), 80, 60, 0, True)
    # icosphere, shared vertices, baked Lambert
    v, i = S.icosphere(1)
    out["ico1"] = (v, i, S.app_transform(0.4), 96, 54, 0, False)
    out["ico1_depth"] = (v, i, S.app_transform(0.4), 96, 54, 0, True)
    out["ico1_vertices"] = (v, i, S.app_transform(0.4), 96, 54, 2, False)
    out["ico1_line"] = (v, i[: (i.size // 2) * 2], S.app_transform(0.4), 32, 18, 1, False)
    # random soup: overlap, off-screen corners, colours outside [0,1], negative w
    n = 60
    xyz = rng.uniform(-1.4, 1.4, size=(3 * n, 3)).astype(np.float32)
    xyz[:, 2] = rng.uniform(-0.3, 1.0, size=3 * n)
    rgb = rng.uniform(-0.3, 1.3, size=(3 * n, 3)).astype(np.float32)
    v = S.pack_vertices(xyz, rgb)
    i = np.arange(3 * n)
    persp = np.array([[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, 1, 0], [0, 0, 1, 1]], np.float32)     # App.swift:176-181
    out["soup60"] = (v, i, I4, 64, 48, 0, False)
    out["soup60_depth"] = (v, i, I4, 64, 48, 0, True)
    out["soup60_persp"] = (v, i, persp, 64, 48, 0, False)
    # integer-degenerate and tie cases: sub-pixel triangles, equal-y rows, equal depths, duplicate triangles
    small = (rng.uniform(-0.9, 0.9, size=(40, 1, 3)) + rng.uniform(-0.03, 0.03, size=(40, 3, 3))).reshape(-1, 3).astype(np.float32)
    out["tiny40"] = (S.pack_vertices(small, rng.uniform(0, 1, size=(120, 3)).astype(np.float32)), np.arange(120), I4, 48, 48, 0, False)
    flat = tri([(-0.5, 0.25), (0.1, 0.25), (0.6, 0.25)])
    out["flat_row"] = (flat, np.arange(3), I4, 40, 40, 0, False)
    two = np.concatenate([tri([(0, 1), (1, -1), (-1, -1)], z=(0.5, 0.5, 0.5)),
                          tri([(0, -1), (1, 1), (-1, 1)], z=(0.5, 0.5, 0.5), rgb=[[1, 1, 0], [0, 1, 1], [1, 0, 1]])])
    out["tie_depth"] = (two, np.arange(6), I4, 40, 40, 0, True)
    out["tie_order"] = (two, np.arange(6), I4, 40, 40, 0, False)
    # domain rule: one triangle with w = 0 (inf screen coordinates) is skipped and counted
    bad = np.concatenate([tri([(0, 1), (1, 0), (-1, 0)], z=(-1, 0, 0)), tri([(0, 0.5), (0.5, -0.5), (-0.5, -0.5)])])
    out["domain_skip"] = (bad, np.arange(6), persp, 32, 32, 0, False)
    return out


def main():
    store = {}
    names = []
    for name, (v, i, m, W, H, prim, depth) in cases().items():
        v = np.ascontiguousarray(v, np.float32)
        i = np.ascontiguousarray(i, np.int64)
        m = np.ascontiguousarray(m, np.float32)
        flags = O.LITERAL | (O.DEPTH_LESS if depth else 0)
        rc, c_cpp, d_cpp, st = O.render(v, i, m, W, H, primitive_type=prim, flags=flags)
        c_py, d_py, skipped = P.render(v, i, m, W, H, primitive_type=prim, depth_less=depth)
        assert rc in (O.OK, O.ERR_DOMAIN), (name, rc)
        assert st["domain_skipped"] == skipped, (name, st, skipped)
        bad = (c_cpp != c_py).any(-1)
        assert not bad.any(), f"{name}: C++ and Python restatements disagree on {int(bad.sum())} colour pixels, first {np.argwhere(bad)[0]}"
        assert np.array_equal(d_cpp.view(np.uint32), d_py.view(np.uint32)), f"{name}: depth disagrees"
        names.append(name)
        store[f"{name}/vertices"] = v
        store[f"{name}/indices"] = i
        store[f"{name}/transform"] = m
        store[f"{name}/meta"] = np.array([W, H, prim, int(depth), skipped], np.int64)
        store[f"{name}/color"] = c_cpp
        store[f"{name}/depth"] = d_cpp
        print(f"{name:16s} {W}x{H} prims={i.size // (2 if prim == 1 else 3):4d} covered={(c_cpp[..., 3] > 0).sum():5d} skipped={skipped}  both restatements agree")
    store["names"] = np.array(names)
    np.savez_compressed(OUT, **store)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
