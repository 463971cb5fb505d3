"""Caller-side scene code: procedural meshes, baked vertex colours and transform composition.

This is what sits ABOVE render(renderPass:) in the reference (App.swift:112-134 mesh import,
App.swift:163-183 transform composition).  It is host-side numpy: meshes are built once and uploaded
with CudaRenderer.set_mesh; nothing here is on the timed path.

Vertex rows are laid out like Swift's `Vertex` (Renderer.swift:154-157): x y z pad r g b pad (32 bytes).
Indices are int64 like Swift's `[Int]` (Renderer.swift:196).
"""
from __future__ import annotations

import numpy as np

# ---------------------------------------------------------------------------------------------
# transforms (App.swift:169-183)
# ---------------------------------------------------------------------------------------------

#: projection literal of App.swift:176-181 (rows): w = z + 1
PROJECTION = np.array([[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, 1, 0], [0, 0, 1, 1]], dtype=np.float64)


def quat(angle: float, axis) -> np.ndarray:
    """simd_quatf(angle:axis:) -> (x, y, z, w)."""
    a = np.asarray(axis, dtype=np.float64)
    s = np.sin(angle * 0.5)
    return np.array([a[0] * s, a[1] * s, a[2] * s, np.cos(angle * 0.5)], dtype=np.float64)


def quat_mul(p, q) -> np.ndarray:
    px, py, pz, pw = p
    qx, qy, qz, qw = q
    return np.array([
        pw * qx + px * qw + py * qz - pz * qy,
        pw * qy - px * qz + py * qw + pz * qx,
        pw * qz + px * qy - py * qx + pz * qw,
        pw * qw - px * qx - py * qy - pz * qz], dtype=np.float64)


def quat_matrix(q) -> np.ndarray:
    x, y, z, w = q
    return np.array([
        [1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w), 0],
        [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w), 0],
        [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y), 0],
        [0, 0, 0, 1]], dtype=np.float64)


def trs(translation=(0, 0, 0), rotation=(0, 0, 0, 1), scale=(1, 1, 1)) -> np.ndarray:
    """RealityKit Transform.matrix = T * R * S (App.swift:169-174)."""
    t = np.eye(4)
    t[:3, 3] = translation
    s = np.diag([scale[0], scale[1], scale[2], 1.0])
    return t @ quat_matrix(rotation) @ s


def app_transform(time: float) -> np.ndarray:
    """The matrix App.swift:169-183 builds for animation time `time` (float32, M @ v convention)."""
    n = np.array([1.0, 1.0, 0.0]) / np.sqrt(2.0)
    rot = quat_mul(quat(time, n), quat(time * 0.5, (0.0, 0.0, 1.0)))
    m = PROJECTION @ trs(translation=(0, 0, 1), rotation=rot, scale=(2, 2, 2))
    return m.astype(np.float32)


def view_transform(direction, roll: float = 0.0) -> np.ndarray:
    """Camera looking at the mesh from `direction`: rotate the mesh so `direction` faces the eye, then the
    App.swift placement (scale 2, translate z=1, projection)."""
    d = np.asarray(direction, dtype=np.float64)
    d = d / np.linalg.norm(d)
    target = np.array([0.0, 0.0, -1.0])
    axis = np.cross(d, target)
    s = np.linalg.norm(axis)
    c = float(np.dot(d, target))
    if s < 1e-12:
        q = quat(0.0 if c > 0 else np.pi, (1.0, 0.0, 0.0))
    else:
        q = quat(np.arctan2(s, c), axis / s)
    q = quat_mul(quat(roll, (0.0, 0.0, 1.0)), q)
    m = PROJECTION @ trs(translation=(0, 0, 1), rotation=q, scale=(2, 2, 2))
    return m.astype(np.float32)


def fibonacci_views(n: int) -> np.ndarray:
    """n camera matrices on a Fibonacci sphere (config 5).  Returns (n,4,4) float32, M @ v convention."""
    out = np.empty((n, 4, 4), dtype=np.float32)
    golden = np.pi * (3.0 - np.sqrt(5.0))
    for i in range(n):
        z = 1.0 - 2.0 * (i + 0.5) / n
        r = np.sqrt(max(0.0, 1.0 - z * z))
        th = golden * i
        out[i] = view_transform((r * np.cos(th), r * np.sin(th), z), roll=0.37 * i)
    return out


def soup_transform(k: int) -> np.ndarray:
    """Per-frame matrix for the triangle soup: projection * small in-plane rotation (frame k)."""
    th = 0.002 * k
    rz = np.array([[np.cos(th), -np.sin(th), 0, 0], [np.sin(th), np.cos(th), 0, 0], [0, 0, 1, 0], [0, 0, 0, 1]])
    return (PROJECTION @ rz).astype(np.float32)


# ---------------------------------------------------------------------------------------------
# meshes
# ---------------------------------------------------------------------------------------------

def pack_vertices(xyz: np.ndarray, rgb: np.ndarray) -> np.ndarray:
    v = np.zeros((xyz.shape[0], 8), dtype=np.float32)
    v[:, 0:3] = xyz
    v[:, 4:7] = rgb
    return v


def cube(side: float = 0.5):
    """24-vertex cube, per-face colour = abs(face normal) (cf. App.swift:125,133).  12 triangles."""
    h = side / 2
    xyz, rgb, idx = [], [], []
    for axis in range(3):
        for sign in (-1.0, 1.0):
            u, v = (axis + 1) % 3, (axis + 2) % 3
            base = len(xyz)
            for du, dv in ((-1, -1), (1, -1), (1, 1), (-1, 1)):
                p = [0.0, 0.0, 0.0]
                p[axis] = sign * h
                p[u] = du * h
                p[v] = dv * h
                xyz.append(p)
                n = [0.0, 0.0, 0.0]
                n[axis] = 1.0
                rgb.append(n)
            quad = (0, 1, 2, 0, 2, 3) if sign > 0 else (0, 2, 1, 0, 3, 2)
            idx += [base + q for q in quad]
    return pack_vertices(np.array(xyz, np.float32), np.array(rgb, np.float32)), np.array(idx, dtype=np.int64)


def lambert(normals: np.ndarray, albedo: float = 0.8, light=(1.0, 1.0, -1.0)) -> np.ndarray:
    """Bake albedo * max(0, n.l) into the vertex colour (the reference has no lighting: Renderer.swift:263-268)."""
    l = np.asarray(light, dtype=np.float64)
    l = l / np.linalg.norm(l)
    k = albedo * np.maximum(0.0, normals.astype(np.float64) @ l)
    return np.repeat(k[:, None], 3, axis=1).astype(np.float32)


def icosphere(subdiv: int = 6, radius: float = 0.4, shade: str = "lambert"):
    """Icosphere with shared vertices: 20*4^s triangles, 10*4^s+2 vertices."""
    t = (1.0 + np.sqrt(5.0)) / 2.0
    verts = np.array([[-1, t, 0], [1, t, 0], [-1, -t, 0], [1, -t, 0], [0, -1, t], [0, 1, t], [0, -1, -t], [0, 1, -t],
                      [t, 0, -1], [t, 0, 1], [-t, 0, -1], [-t, 0, 1]], dtype=np.float64)
    verts /= np.linalg.norm(verts, axis=1, keepdims=True)
    faces = np.array([[0, 11, 5], [0, 5, 1], [0, 1, 7], [0, 7, 10], [0, 10, 11], [1, 5, 9], [5, 11, 4], [11, 10, 2],
                      [10, 7, 6], [7, 1, 8], [3, 9, 4], [3, 4, 2], [3, 2, 6], [3, 6, 8], [3, 8, 9], [4, 9, 5],
                      [2, 4, 11], [6, 2, 10], [8, 6, 7], [9, 8, 1]], dtype=np.int64)
    for _ in range(subdiv):
        nv = verts.shape[0]
        e = np.concatenate([faces[:, [0, 1]], faces[:, [1, 2]], faces[:, [2, 0]]], axis=0)
        e_sorted = np.sort(e, axis=1)
        key = e_sorted[:, 0] * nv + e_sorted[:, 1]
        uniq, inv = np.unique(key, return_inverse=True)
        a, b = uniq // nv, uniq % nv
        mid = verts[a] + verts[b]
        mid /= np.linalg.norm(mid, axis=1, keepdims=True)
        verts = np.concatenate([verts, mid], axis=0)
        nf = faces.shape[0]
        m01, m12, m20 = nv + inv[:nf], nv + inv[nf:2 * nf], nv + inv[2 * nf:]
        v0, v1, v2 = faces[:, 0], faces[:, 1], faces[:, 2]
        faces = np.concatenate([np.stack([v0, m01, m20], 1), np.stack([v1, m12, m01], 1),
                                np.stack([v2, m20, m12], 1), np.stack([m01, m12, m20], 1)], axis=0)
    rgb = lambert(verts) if shade == "lambert" else np.abs(verts).astype(np.float32)
    return pack_vertices((verts * radius).astype(np.float32), rgb), faces.reshape(-1).astype(np.int64)


def cube_sphere(n: int, radius: float = 0.4, shade: str = "normal"):
    """Cube with n x n quads per face, welded and pushed onto the sphere: 12 n^2 triangles, 6 n^2 + 2 vertices."""
    g = np.arange(n + 1, dtype=np.int64)
    uu, vv = np.meshgrid(g, g, indexing="ij")
    uu, vv = uu.reshape(-1), vv.reshape(-1)
    pts, tris = [], []
    for axis in range(3):
        for side in (0, n):
            p = np.empty((uu.size, 3), dtype=np.int64)
            p[:, axis] = side
            p[:, (axis + 1) % 3] = uu
            p[:, (axis + 2) % 3] = vv
            base = len(pts) * (n + 1) * (n + 1)
            pts.append(p)
            i, j = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
            i, j = i.reshape(-1), j.reshape(-1)
            q00 = base + i * (n + 1) + j
            q10 = q00 + (n + 1)
            q01 = q00 + 1
            q11 = q10 + 1
            if side == n:
                t = np.stack([q00, q10, q11, q00, q11, q01], axis=1)
            else:
                t = np.stack([q00, q11, q10, q00, q01, q11], axis=1)
            tris.append(t.reshape(-1))
    pts = np.concatenate(pts, axis=0)
    tris = np.concatenate(tris, axis=0)
    key = (pts[:, 0] * (n + 1) + pts[:, 1]) * (n + 1) + pts[:, 2]
    uniq, first, inv = np.unique(key, return_index=True, return_inverse=True)
    p = pts[first].astype(np.float64) / n * 2.0 - 1.0
    p /= np.linalg.norm(p, axis=1, keepdims=True)
    rgb = lambert(p) if shade == "lambert" else np.abs(p).astype(np.float32)
    return pack_vertices((p * radius).astype(np.float32), rgb), inv[tris].astype(np.int64)


def triangle_soup(count: int, width: int, height: int, seed: int = 0x5EED0003,
                  area_px=(1.0, 1600.0), z_range=(0.2, 1.8)):
    """`count` independent triangles (3*count unshared vertices) for use with PROJECTION (w = z + 1):
    NDC centre uniform in [-1,1]^2, projected area log-uniform in `area_px` pixels at width x height,
    z uniform in `z_range` with a small per-corner offset, colours uniform."""
    rng = np.random.Generator(np.random.PCG64(seed))
    c = rng.uniform(-1.0, 1.0, size=(count, 1, 2))
    area = np.exp(rng.uniform(np.log(area_px[0]), np.log(area_px[1]), size=count))
    shape = rng.standard_normal(size=(count, 3, 2))
    shape -= shape.mean(axis=1, keepdims=True)
    e1, e2 = shape[:, 1] - shape[:, 0], shape[:, 2] - shape[:, 0]
    a0 = 0.5 * np.abs(e1[:, 0] * e2[:, 1] - e1[:, 1] * e2[:, 0])
    a0 = np.maximum(a0, 1e-3)
    shape *= np.sqrt(area / a0)[:, None, None]          # pixel offsets with the requested area
    shape = np.clip(shape, -200.0, 200.0)               # bound sliver extents
    ndc = c + shape * np.array([2.0 / width, 2.0 / height])
    z = rng.uniform(z_range[0], z_range[1], size=(count, 1)) + rng.uniform(-0.02, 0.02, size=(count, 3))
    w = z + 1.0
    xyz = np.empty((count, 3, 3), dtype=np.float64)
    xyz[:, :, 0] = ndc[:, :, 0] * w
    xyz[:, :, 1] = ndc[:, :, 1] * w
    xyz[:, :, 2] = z
    rgb = rng.uniform(0.0, 1.0, size=(count, 3, 3))
    verts = pack_vertices(xyz.reshape(-1, 3).astype(np.float32), rgb.reshape(-1, 3).astype(np.float32))
    return verts, np.arange(3 * count, dtype=np.int64)


def algorithmic_bytes(ntris: int, nverts: int, width: int, height: int) -> int:
    """each index once as int32, each vertex once as xyz+rgb, each pixel's colour and depth once."""
    return 12 * ntris + 24 * nverts + 8 * width * height


# ---------------------------------------------------------------------------------------------
# headless image output (replaces the MTKView blit + present of App.swift:103-105)
# ---------------------------------------------------------------------------------------------
def write_ppm(path: str, bgra: np.ndarray) -> None:
    """Binary PPM (P6) of a (H, W, 4) BGRA8 image as the renderer writes it (Pixel = b,g,r,a, Renderer.swift:44-49)."""
    img = np.asarray(bgra, dtype=np.uint8)
    h, w = img.shape[:2]
    with open(path, "wb") as f:
        f.write(b"P6\n%d %d\n255\n" % (w, h))
        f.write(np.ascontiguousarray(img[..., 2::-1]).tobytes())


def write_png(path: str, bgra: np.ndarray) -> None:
    """RGBA PNG (zlib only, no imaging dependency) of a (H, W, 4) BGRA8 image."""
    import struct
    import zlib
    img = np.asarray(bgra, dtype=np.uint8)
    h, w = img.shape[:2]
    rgba = np.ascontiguousarray(img[..., [2, 1, 0, 3]])
    raw = np.concatenate([np.zeros((h, 1), np.uint8), rgba.reshape(h, w * 4)], axis=1).tobytes()   # filter 0 per row

    def chunk(tag: bytes, data: bytes) -> bytes:
        return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)

    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 6, 0, 0, 0))
                + chunk(b"IDAT", zlib.compress(raw, 6)) + chunk(b"IEND", b""))
