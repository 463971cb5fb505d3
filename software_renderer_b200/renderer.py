"""Host-side mirror of the reference's render interface over the C ABI of libsr_cuda.so.

The reference's seam is `Renderer.render(renderPass:)` (Renderer.swift:204-230; the Metal twin is
GpuRenderer.swift:35).  `CudaRenderer.render(render_pass)` keeps that shape — same `RenderPass` fields
(Renderer.swift:191-200), same `PrimitiveType` cases (Renderer.swift:174-189), caller-owned colour/depth images
written in place (App.swift:80-101) — and forwards to the CUDA library through ctypes.  The Swift binding of the
same ABI is under swift/ (it cannot be compiled in this image: no Swift toolchain).

There is no CPU fallback: if libsr_cuda.so is missing, or no CUDA device is present, construction raises.
"""
from __future__ import annotations

import ctypes
import enum
import os
from dataclasses import dataclass, field

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SR_LIB") or os.path.join(_HERE, "libsr_cuda.so")   # SR_LIB: a debug build (e.g. -DSR_TIMELINE)
#: the -DSR_SIMD_FMA build: Apple's simd_muladd fused, as arm64 compiles it; CudaRenderer(simd_fma=True)
LIB_PATH_FMA = os.path.join(_HERE, "libsr_cuda_fma.so")

SR_OK, SR_ERR_ARG, SR_ERR_INDEX, SR_ERR_DOMAIN, SR_ERR_CUDA, SR_ERR_NOMESH, SR_ERR_COMM = 0, -1, -2, -3, -4, -5, -6
FLAG_DEPTH_LESS = 1 << 0
FLAG_NO_DEPTH_OUT = 1 << 1
FLAG_DEPTH_UNTOUCHED = 1 << 2

#: every symbol include/sr_cuda.h declares
ABI_SYMBOLS = (
    "sr_version", "sr_create", "sr_destroy", "sr_last_error", "sr_set_mesh", "sr_render", "sr_render_pass",
    "sr_render_batch", "sr_render_device", "sr_sync", "sr_get_stats", "sr_get_kernel_times", "sr_set_profiling",
    "sr_draw2d", "sr_set_band", "sr_mesh_generation", "sr_host_alloc", "sr_host_free", "sr_host_register",
    "sr_host_unregister", "sr_host_fill_depth", "sr_comm_unique_id", "sr_comm_init", "sr_comm_destroy", "sr_gather", "sr_gather_fence",
    "sr_gather_sync", "sr_points_no_alloc", "sr_points_no_alloc_last_error",
)


class PrimitiveType(enum.IntEnum):
    """Renderer.swift:174-189."""
    triangle = 0
    line = 1
    vertices = 2

    @property
    def vertices_count(self) -> int:
        return 2 if self is PrimitiveType.line else 3


class Cmd2D(ctypes.Structure):
    """sr_cmd2d: one command of the legacy 2-D immediate-mode API (Renderer.swift:376-465)."""
    _fields_ = [("kind", ctypes.c_int32), ("p", ctypes.c_int32 * 6), ("color", ctypes.c_uint8 * 4)]

    RECT, CIRCLE, LINE, TRIANGLE = 0, 1, 2, 3

    @classmethod
    def make(cls, kind, params, color=(0, 0, 0, 0)):
        c = cls()
        c.kind = kind
        for k, v in enumerate(params):
            c.p[k] = int(v)
        for k, v in enumerate(color):
            c.color[k] = int(v)
        return c

    @classmethod
    def rect(cls, x, y, w, h, color):            # draw(rect:with:in:)  — note: w drives y and h drives x (R:378-379)
        return cls.make(cls.RECT, (x, y, w, h), color)

    @classmethod
    def circle(cls, x, y, r):                    # draw(circle:with:in:) — the colour argument is ignored by the reference
        return cls.make(cls.CIRCLE, (x, y, r))

    @classmethod
    def line(cls, x0, y0, x1, y1, color):        # draw(line:with:in:)
        return cls.make(cls.LINE, (x0, y0, x1, y1), color)

    @classmethod
    def triangle(cls, a, b, c):                  # draw(triangle:with:in:) — colour = barycentric weights
        return cls.make(cls.TRIANGLE, (a[0], a[1], b[0], b[1], c[0], c[1]))


def pack_cmds(cmds):
    arr = (Cmd2D * max(len(cmds), 1))()
    for k, c in enumerate(cmds):
        arr[k] = c
    return arr


class SrStats(ctypes.Structure):
    _fields_ = [
        ("primitives", ctypes.c_uint64), ("domain_skipped", ctypes.c_uint64), ("small_tris", ctypes.c_uint64),
        ("binned_tris", ctypes.c_uint64), ("tile_pairs", ctypes.c_uint64), ("bin_regrows", ctypes.c_uint64),
        ("pairs_consumed", ctypes.c_uint64), ("batches_run", ctypes.c_uint64), ("depth_culled", ctypes.c_uint64), ("nobin_reruns", ctypes.c_uint64),
        ("depth_two_pass_frames", ctypes.c_uint64), ("back_survivors", ctypes.c_uint64),
        ("gpu_launches", ctypes.c_uint64), ("gpu_ms", ctypes.c_float),
    ]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class SrError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"libsr_cuda error {code}: {message}")
        self.code = code


class DomainError(SrError):
    """>= 1 primitive was skipped because a corner left the valid domain (the reference traps, Renderer.swift:251).
    The frame was completed; the images hold everything else."""


_libs = {}


def load_library(simd_fma: bool = False):
    """Loads libsr_cuda.so (built in-tree by __graft_entry__.build()), or its SR_SIMD_FMA twin.  Raises if it is missing."""
    path = LIB_PATH_FMA if simd_fma else LIB_PATH
    if path in _libs:
        return _libs[path]
    if not os.path.exists(path):
        raise FileNotFoundError(
            f"{path} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU fallback)")
    L = ctypes.CDLL(path)
    vp, sz, i32, u32 = ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_uint32
    L.sr_version.restype = i32
    L.sr_version.argtypes = []
    L.sr_create.restype = i32
    L.sr_create.argtypes = [ctypes.POINTER(vp), i32]
    L.sr_destroy.restype = None
    L.sr_destroy.argtypes = [vp]
    L.sr_last_error.restype = ctypes.c_char_p
    L.sr_last_error.argtypes = [vp]
    L.sr_set_mesh.restype = i32
    L.sr_set_mesh.argtypes = [vp, vp, sz, vp, sz]
    L.sr_render.restype = i32
    L.sr_render.argtypes = [vp, vp, i32, i32, i32, vp, vp, u32]
    L.sr_render_pass.restype = i32
    L.sr_render_pass.argtypes = [vp, vp, sz, vp, sz, vp, i32, i32, i32, vp, vp, u32]
    L.sr_render_batch.restype = i32
    L.sr_render_batch.argtypes = [vp, vp, sz, i32, i32, i32, vp, vp, u32]
    L.sr_render_device.restype = i32
    L.sr_render_device.argtypes = [vp, vp, sz, i32, i32, i32, vp, vp, u32, vp]
    L.sr_sync.restype = i32
    L.sr_sync.argtypes = [vp]
    L.sr_get_stats.restype = i32
    L.sr_get_stats.argtypes = [vp, ctypes.POINTER(SrStats)]
    L.sr_get_kernel_times.restype = i32
    L.sr_get_kernel_times.argtypes = [vp, ctypes.POINTER(ctypes.c_char_p), ctypes.POINTER(ctypes.c_float), i32]
    L.sr_set_profiling.restype = i32
    L.sr_set_profiling.argtypes = [vp, i32]
    L.sr_draw2d.restype = i32
    L.sr_draw2d.argtypes = [vp, vp, sz, i32, i32, vp]
    L.sr_set_band.restype = i32
    L.sr_set_band.argtypes = [vp, i32, i32]
    L.sr_mesh_generation.restype = ctypes.c_uint64
    L.sr_mesh_generation.argtypes = [vp]
    L.sr_host_alloc.restype = i32
    L.sr_host_alloc.argtypes = [vp, sz, ctypes.POINTER(vp)]
    L.sr_host_free.restype = i32
    L.sr_host_free.argtypes = [vp, vp]
    L.sr_host_register.restype = i32
    L.sr_host_register.argtypes = [vp, vp, sz]
    L.sr_host_unregister.restype = i32
    L.sr_host_unregister.argtypes = [vp, vp]
    L.sr_host_fill_depth.restype = i32
    L.sr_host_fill_depth.argtypes = [vp, vp, sz]
    L.sr_points_no_alloc.restype = i32
    L.sr_points_no_alloc.argtypes = [i32, vp, sz, vp, sz, vp, i32, i32, vp, vp, vp]
    L.sr_points_no_alloc_last_error.restype = ctypes.c_char_p
    L.sr_points_no_alloc_last_error.argtypes = []
    L.sr_comm_unique_id.restype = i32
    L.sr_comm_unique_id.argtypes = [vp]
    L.sr_comm_init.restype = i32
    L.sr_comm_init.argtypes = [vp, vp, i32, i32]
    L.sr_comm_destroy.restype = i32
    L.sr_comm_destroy.argtypes = [vp]
    L.sr_gather.restype = i32
    L.sr_gather.argtypes = [vp, vp, vp, ctypes.POINTER(sz), i32, i32, vp]
    L.sr_gather_fence.restype = i32
    L.sr_gather_fence.argtypes = [vp, i32, vp]
    L.sr_gather_sync.restype = i32
    L.sr_gather_sync.argtypes = [vp]
    _libs[path] = L
    return L


@dataclass
class Image:
    """Image<Pixel> (Renderer.swift:8-42): a non-owning view over caller memory, packed rows."""
    array: np.ndarray   # colour: (H, W, 4) uint8 BGRA;  depth: (H, W) float32
    width: int
    height: int

    @staticmethod
    def color(width: int, height: int) -> "Image":
        return Image(np.zeros((height, width, 4), dtype=np.uint8), width, height)

    @staticmethod
    def depth(width: int, height: int) -> "Image":
        return Image(np.zeros((height, width), dtype=np.float32), width, height)


@dataclass
class RenderPass:
    """Renderer.swift:191-200.  `vertices`: (N, 8) float32 rows x y z _ r g b _ (Swift Vertex, Renderer.swift:154-157);
    `indices`: int64; `transform`: (4, 4) float32 in the M @ v sense (handed over column-major)."""
    color_buffer: Image
    depth_buffer: Image
    vertices: np.ndarray
    indices: np.ndarray
    primitive_type: PrimitiveType = PrimitiveType.triangle
    transform: np.ndarray = field(default_factory=lambda: np.eye(4, dtype=np.float32))
    #: None (default) = the reference's contract: the arrays are handed over and uploaded on EVERY call (App.swift:163).
    #: An integer = "static mesh" opt-in: the arrays are uploaded when this number (or the array sizes) differs from the
    #: one uploaded last; the caller bumps it after editing vertices / indices.  Not a field of the reference's RenderPass.
    mesh_version: int | None = None


def _colmajor(transform) -> np.ndarray:
    m = np.asarray(transform, dtype=np.float32)
    if m.ndim == 2:
        return np.ascontiguousarray(m.T).reshape(16)
    return np.ascontiguousarray(np.transpose(m, (0, 2, 1))).reshape(-1, 16)


class CudaRenderer:
    """Drop-in for `Renderer` / `GpuRenderer` (App.swift:148-154) on one B200."""

    def __init__(self, device: int = 0, simd_fma: bool = False):
        self._lib = load_library(simd_fma)
        self._ctx = ctypes.c_void_p()
        rc = self._lib.sr_create(ctypes.byref(self._ctx), int(device))
        if rc != SR_OK:
            self._ctx = ctypes.c_void_p()
            raise SrError(rc, "sr_create failed: no CUDA device / CUDA runtime error (there is no CPU fallback)")
        self.device = device
        self.depth_test = False      # True = run the depth test of Renderer.swift:257-261
        self.depth_untouched = False  # True = host calls pass SR_FLAG_DEPTH_UNTOUCHED (see include/sr_cuda.h)
        self._mesh_key = None        # (mesh_version, nverts, nidx, library generation) of the static-mesh opt-in
        self._band = None            # (y0, y1): render only these rows of each frame (set_band)
        self._host_blocks = {}       # address -> keep-alive of arrays handed out by host_array()

    # -- lifetime ----------------------------------------------------------------------------------------------
    def close(self):
        if getattr(self, "_ctx", None) and self._ctx.value:
            for addr in list(getattr(self, "_host_blocks", {})):
                self._lib.sr_host_free(self._ctx, ctypes.c_void_p(addr))
            self._host_blocks = {}
            self._lib.sr_destroy(self._ctx)
            self._ctx = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # -- helpers -----------------------------------------------------------------------------------------------
    def _check(self, rc: int):
        if rc == SR_OK:
            return
        msg = (self._lib.sr_last_error(self._ctx) or b"").decode()
        if rc == SR_ERR_DOMAIN:
            raise DomainError(rc, f"{self.stats()['domain_skipped']} primitive(s) skipped by the domain rule")
        raise SrError(rc, msg)

    def _flags(self, want_depth: bool, host: bool = False) -> int:
        return ((FLAG_DEPTH_LESS if self.depth_test else 0) | (0 if want_depth else FLAG_NO_DEPTH_OUT) |
                (FLAG_DEPTH_UNTOUCHED if (host and self.depth_untouched) else 0))

    # -- page-locked host memory (sr_host_alloc) ---------------------------------------------------------------
    def host_array(self, shape, dtype) -> np.ndarray:
        """A numpy array in page-locked memory from sr_host_alloc (freed by close()): images and mesh arrays kept here
        are read / written by the GPU's copy engines directly."""
        dt = np.dtype(dtype)
        n = int(np.prod(shape)) * dt.itemsize
        p = ctypes.c_void_p()
        self._check(self._lib.sr_host_alloc(self._ctx, max(n, 1), ctypes.byref(p)))
        buf = (ctypes.c_uint8 * max(n, 1)).from_address(p.value)
        arr = np.frombuffer(buf, dtype=dt, count=int(np.prod(shape))).reshape(shape)
        self._host_blocks[p.value] = buf
        return arr

    def host_register(self, array: np.ndarray):
        """Page-locks memory the caller already owns (sr_host_register); undo with host_unregister."""
        self._check(self._lib.sr_host_register(self._ctx, array.ctypes.data, array.nbytes))

    def host_fill_depth(self, array: np.ndarray):
        """clear(image:with: .infinity) (Renderer.swift:206) on a float32 host array, by the library's host threads."""
        assert array.dtype == np.float32 and array.flags.c_contiguous
        self._check(self._lib.sr_host_fill_depth(self._ctx, array.ctypes.data, array.size))

    def host_unregister(self, array: np.ndarray):
        self._check(self._lib.sr_host_unregister(self._ctx, array.ctypes.data))

    # -- mesh --------------------------------------------------------------------------------------------------
    @staticmethod
    def _mesh_arrays(vertices, indices):
        v = np.ascontiguousarray(vertices, dtype=np.float32)
        if v.ndim != 2 or v.shape[1] != 8:
            raise ValueError("vertices must be (N, 8) float32 rows: x y z _ r g b _")
        return v, np.ascontiguousarray(indices, dtype=np.int64).reshape(-1)

    def set_mesh(self, vertices: np.ndarray, indices: np.ndarray):
        """Explicit upload (sr_set_mesh) for render_views / render_device and for passes with a `mesh_version`."""
        v, idx = self._mesh_arrays(vertices, indices)
        self._check(self._lib.sr_set_mesh(self._ctx, v.ctypes.data, v.shape[0], idx.ctypes.data, idx.shape[0]))
        self._mesh_key = None

    # -- screen-band split of one frame ------------------------------------------------------------------------
    def set_band(self, y0: int = 0, y1: int = 0):
        """Subsequent render calls produce only rows [y0, y1) of each frame, bit-identical to those rows of the whole
        frame; output images hold y1 - y0 rows.  set_band() returns to whole frames."""
        self._check(self._lib.sr_set_band(self._ctx, int(y0), int(y1)))
        self._band = (int(y0), int(y1)) if y1 > 0 else None

    def _rows(self, height: int) -> int:
        return height if self._band is None else self._band[1] - self._band[0]

    # -- render(renderPass:) -----------------------------------------------------------------------------------
    def render(self, render_pass: RenderPass):
        """Renderer.render(renderPass:) (Renderer.swift:204-230): synchronous, writes both images in place
        (with a band set: rows [y0, y1) of both images, the rest is left untouched).
        The mesh arrays travel with the pass and are uploaded on every call, as the reference hands them over
        (App.swift:163; sr_render_pass) - unless the pass opts into a static mesh with `mesh_version` (then they are
        uploaded when the version or the array sizes change; there is no guessing from addresses or object identity)."""
        rp = render_pass
        col, dep = rp.color_buffer.array, rp.depth_buffer.array
        W, H = rp.color_buffer.width, rp.color_buffer.height
        assert col.dtype == np.uint8 and col.flags.c_contiguous and col.size == W * H * 4
        assert dep.dtype == np.float32 and dep.flags.c_contiguous and dep.size == W * H
        m = _colmajor(rp.transform)
        # with a band set only rows [y0, y1) of the pass's images are written: packed rows, so the band is one block
        y0 = 0 if self._band is None else self._band[0]
        cptr, dptr = col.ctypes.data + 4 * W * y0, dep.ctypes.data + 4 * W * y0
        if rp.mesh_version is None:
            v, idx = self._mesh_arrays(rp.vertices, rp.indices)
            self._mesh_key = None
            self._check(self._lib.sr_render_pass(self._ctx, v.ctypes.data, v.shape[0], idx.ctypes.data, idx.shape[0],
                                                 m.ctypes.data, int(rp.primitive_type), W, H, cptr, dptr, self._flags(True, True)))
            return
        key = (int(rp.mesh_version), int(np.shape(rp.vertices)[0]), int(np.size(rp.indices)))
        if self._mesh_key is None or self._mesh_key[:3] != key or self._mesh_key[3] != self._lib.sr_mesh_generation(self._ctx):
            self.set_mesh(rp.vertices, rp.indices)
            self._mesh_key = key + (self._lib.sr_mesh_generation(self._ctx),)
        self._check(self._lib.sr_render(self._ctx, m.ctypes.data, int(rp.primitive_type), W, H, cptr, dptr, self._flags(True, True)))

    def render_views(self, transforms, width: int, height: int, primitive_type=PrimitiveType.triangle,
                     color_out: np.ndarray | None = None, depth_out: np.ndarray | None = None, want_depth=True):
        """`n` independent render passes of the uploaded mesh.  Returns (color[n,H,W,4], depth[n,H,W] or None);
        with a band set (set_band) the images hold the band's rows only."""
        m = _colmajor(np.asarray(transforms, dtype=np.float32).reshape(-1, 4, 4))
        n = m.shape[0]
        rows = self._rows(height)
        color = color_out if color_out is not None else np.empty((n, rows, width, 4), dtype=np.uint8)
        depth = None
        if want_depth:
            depth = depth_out if depth_out is not None else np.empty((n, rows, width), dtype=np.float32)
        assert color.size >= n * rows * width * 4 and (depth is None or depth.size >= n * rows * width)
        rc = self._lib.sr_render_batch(self._ctx, m.ctypes.data, n, int(primitive_type), width, height,
                                       color.ctypes.data, depth.ctypes.data if depth is not None else None,
                                       self._flags(want_depth, True))
        self._check(rc)
        return color, depth

    def render_device(self, transforms, width: int, height: int, color_ptr: int, depth_ptr: int | None,
                      primitive_type=PrimitiveType.triangle, stream: int | None = None, sync: bool = True,
                      colmajor: bool = False):
        """Device-resident variant: `color_ptr` / `depth_ptr` are CUDA device addresses (e.g. torch tensors'
        data_ptr()) for n consecutive images; `stream` is a cudaStream_t handle or None.  colmajor=True: `transforms`
        already is the (n, 16) float32 array CudaRenderer.colmajor() returns."""
        m = transforms if colmajor else _colmajor(np.asarray(transforms, dtype=np.float32).reshape(-1, 4, 4))
        rc = self._lib.sr_render_device(self._ctx, m.ctypes.data, m.shape[0], int(primitive_type), width, height,
                                        ctypes.c_void_p(color_ptr), ctypes.c_void_p(depth_ptr) if depth_ptr else None,
                                        self._flags(depth_ptr is not None), ctypes.c_void_p(stream) if stream else None)
        self._check(rc)
        if sync:
            self.sync()

    @staticmethod
    def colmajor(transforms) -> np.ndarray:
        """(n, 16) float32 column-major copies of (n, 4, 4) / (4, 4) matrices, the layout the C ABI takes: convert once,
        then enqueue with render_device(..., colmajor=True) from a loop that must stay ahead of the GPU."""
        return _colmajor(np.asarray(transforms, dtype=np.float32).reshape(-1, 4, 4))

    def sync(self):
        self._check(self._lib.sr_sync(self._ctx))

    # -- composite gather over NCCL (sr_gather) -----------------------------------------------
    @staticmethod
    def comm_unique_id(simd_fma: bool = False) -> bytes:
        """The 128-byte NCCL id the root rank creates and hands to the others (sr_comm_unique_id)."""
        buf = ctypes.create_string_buffer(128)
        rc = load_library(simd_fma).sr_comm_unique_id(buf)
        if rc != SR_OK:
            raise SrError(rc, "sr_comm_unique_id failed: libnccl.so.2 could not be opened")
        return buf.raw

    def comm_init(self, unique_id: bytes, rank: int, world: int):
        assert len(unique_id) == 128
        self._check(self._lib.sr_comm_init(self._ctx, ctypes.c_char_p(unique_id), int(rank), int(world)))
        self.comm_rank, self.comm_world = int(rank), int(world)

    def comm_destroy(self):
        self._lib.sr_comm_destroy(self._ctx)

    def gather(self, send_ptr: int, recv_ptr: int | None, bytes_per_rank, root: int = 0, slot: int = 0,
               after_stream: int | None = None):
        """Asynchronous gather of every rank's block to `root` (device pointers), ordered after the work this context
        has enqueued on `after_stream` so far; runs on the context's gather stream."""
        arr = (ctypes.c_size_t * len(bytes_per_rank))(*[int(b) for b in bytes_per_rank])
        self._check(self._lib.sr_gather(self._ctx, ctypes.c_void_p(send_ptr) if send_ptr else None,
                                        ctypes.c_void_p(recv_ptr) if recv_ptr else None, arr, int(root), int(slot),
                                        ctypes.c_void_p(after_stream) if after_stream else None))

    def gather_fence(self, slot: int, stream: int | None = None):
        self._check(self._lib.sr_gather_fence(self._ctx, int(slot), ctypes.c_void_p(stream) if stream else None))

    def gather_sync(self):
        self._check(self._lib.sr_gather_sync(self._ctx))

    def draw2d(self, cmds, image: np.ndarray):
        """Runs legacy 2-D commands (Cmd2D.rect/circle/line/triangle) in order on a (H, W, 4) BGRA8 image, in place."""
        assert image.dtype == np.uint8 and image.ndim == 3 and image.shape[2] == 4 and image.flags.c_contiguous
        arr = pack_cmds(cmds)
        self._check(self._lib.sr_draw2d(self._ctx, arr, len(cmds), image.shape[1], image.shape[0], image.ctypes.data))
        return image

    # -- introspection -----------------------------------------------------------------------------------------
    def stats(self) -> dict:
        s = SrStats()
        self._lib.sr_get_stats(self._ctx, ctypes.byref(s))
        return s.as_dict()

    def set_profiling(self, enabled: bool):
        self._lib.sr_set_profiling(self._ctx, 1 if enabled else 0)

    def kernel_times(self) -> dict:
        names = (ctypes.c_char_p * 32)()
        ms = (ctypes.c_float * 32)()
        n = self._lib.sr_get_kernel_times(self._ctx, names, ms, 32)
        return {names[i].decode(): float(ms[i]) for i in range(n)}
