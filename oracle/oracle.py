"""ctypes binding of the CPU oracle (oracle/libsr_oracle.so).

TEST INFRASTRUCTURE ONLY (see oracle/sr_oracle.h): imported by tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py, never by the product package.
The oracle restates /root/reference/renderer/Renderer.swift:204-287, 467-494, 88-100, 117-128.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libsr_oracle.so")
_LIB_PATH_FMA = os.path.join(_HERE, "libsr_oracle_fma.so")     # -DSR_SIMD_FMA: Apple's simd_muladd fused (arm64)

TRIANGLE, LINE, VERTICES = 0, 1, 2
DEPTH_LESS = 1 << 0
LITERAL = 1 << 8
OK, ERR_ARG, ERR_INDEX, ERR_DOMAIN = 0, -1, -2, -3


class Stats(ctypes.Structure):
    _fields_ = [
        ("primitives", ctypes.c_uint64),
        ("domain_skipped", ctypes.c_uint64),
        ("span_pixels", ctypes.c_uint64),
        ("rejected_pixels", ctypes.c_uint64),
        ("depth_passed", ctypes.c_uint64),
    ]

    def as_dict(self):
        return {k: int(getattr(self, k)) for k, _ in self._fields_}


def build(force: bool = False) -> str:
    """Compile the oracle with its own Makefile (g++ -O2 -ffp-contract=off)."""
    src = os.path.join(_HERE, "sr_oracle.cpp")
    newest = max(os.path.getmtime(src), os.path.getmtime(os.path.join(_HERE, "sr_oracle.h")))
    stale = any((not os.path.exists(q)) or os.path.getmtime(q) < newest for q in (_LIB_PATH, _LIB_PATH_FMA))
    if force or stale:
        subprocess.run(["make", "-C", _HERE, "clean", "all"], check=True, stdout=subprocess.DEVNULL)
    return _LIB_PATH


_libs = {}


def lib(simd_fma: bool = False):
    """The oracle library: the default (every * and + rounded separately) or the SR_SIMD_FMA variant."""
    if simd_fma not in _libs:
        build()
        L = ctypes.CDLL(_LIB_PATH_FMA if simd_fma else _LIB_PATH)
        L.sr_oracle_render.restype = ctypes.c_int
        L.sr_oracle_render.argtypes = [
            ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_int,
            ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint32, ctypes.POINTER(Stats)]
        L.sr_oracle_transform_vertex.restype = None
        L.sr_oracle_transform_vertex.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
        L.sr_oracle_interpolate.restype = ctypes.c_int64
        L.sr_oracle_interpolate.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int64]
        L.sr_oracle_weights.restype = None
        L.sr_oracle_weights.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_int64, ctypes.c_void_p]
        L.sr_oracle_pack_pixel.restype = None
        L.sr_oracle_pack_pixel.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        L.sr_oracle_draw2d.restype = ctypes.c_int
        L.sr_oracle_draw2d.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
        _libs[simd_fma] = L
    return _libs[simd_fma]


def _vertices32(vertices: np.ndarray) -> np.ndarray:
    """(N,8) float32 rows laid out like Swift's Vertex (xyz pad rgb pad)."""
    v = np.ascontiguousarray(vertices, dtype=np.float32)
    assert v.ndim == 2 and v.shape[1] == 8, "vertices must be (N, 8) float32: x y z _ r g b _"
    return v


def render(vertices, indices, transform, width, height, primitive_type=TRIANGLE, flags=0,
           color_out=None, depth_out=None, simd_fma=False):
    """Returns (rc, color[H,W,4] uint8 BGRA, depth[H,W] float32, stats dict).

    `transform` is a (4,4) numpy matrix in the usual row-major maths sense (M @ v); it is handed to the
    oracle column-major like matrix_float4x4.
    """
    v = _vertices32(vertices)
    idx = np.ascontiguousarray(indices, dtype=np.int64).reshape(-1)
    m = np.ascontiguousarray(np.asarray(transform, dtype=np.float32).T)  # column-major bytes
    color = color_out if color_out is not None else np.empty((height, width, 4), dtype=np.uint8)
    depth = depth_out if depth_out is not None else np.empty((height, width), dtype=np.float32)
    st = Stats()
    rc = lib(simd_fma).sr_oracle_render(v.ctypes.data, v.shape[0], idx.ctypes.data, idx.shape[0], m.ctypes.data,
                                int(primitive_type), int(width), int(height), color.ctypes.data, depth.ctypes.data,
                                int(flags), ctypes.byref(st))
    return rc, color, depth, st.as_dict()


def transform_vertex(transform, xyz, width, height):
    m = np.ascontiguousarray(np.asarray(transform, dtype=np.float32).T)
    p = np.ascontiguousarray(xyz, dtype=np.float32)
    out = np.empty(3, dtype=np.float32)
    lib().sr_oracle_transform_vertex(m.ctypes.data, p.ctypes.data, width, height, out.ctypes.data)
    return out


def interpolate(points, t):
    p = np.ascontiguousarray(points, dtype=np.int64).reshape(-1, 2)
    return int(lib().sr_oracle_interpolate(p.ctypes.data, p.shape[0], int(t)))


def weights(abc, x, y):
    p = np.ascontiguousarray(abc, dtype=np.int64).reshape(3, 2)
    out = np.empty(3, dtype=np.float32)
    lib().sr_oracle_weights(p.ctypes.data, int(x), int(y), out.ctypes.data)
    return out


def pack_pixel(rgb):
    c = np.ascontiguousarray(rgb, dtype=np.float32)
    out = np.empty(4, dtype=np.uint8)
    lib().sr_oracle_pack_pixel(c.ctypes.data, out.ctypes.data)
    return out


def draw2d(cmd_array, ncmds, image):
    """Legacy 2-D API (Renderer.swift:376-465) on a (H, W, 4) BGRA8 image, in place.  `cmd_array` is a ctypes array
    with the sr_oracle_cmd2d layout (software_renderer_b200.renderer.pack_cmds builds one)."""
    assert image.dtype == np.uint8 and image.flags.c_contiguous
    return lib().sr_oracle_draw2d(cmd_array, ncmds, image.shape[1], image.shape[0], image.ctypes.data)
