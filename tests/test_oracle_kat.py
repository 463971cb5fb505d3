"""CPU tests of the oracle: the two hand-derived known-answer vectors of SURVEY.md §4 (the reference ships no
tests or golden vectors), literal-vs-hoisted identity, and the quirks of the span rule (Renderer.swift:467-494)."""
import numpy as np
import pytest

from oracle import oracle as O
from software_renderer_b200 import scene as S

I4 = np.eye(4, dtype=np.float32)


def tri(xy, colors=None):
    xyz = np.array([[x, y, 0.0] for x, y in xy], np.float32)
    rgb = np.eye(3, dtype=np.float32) if colors is None else np.asarray(colors, np.float32)
    return S.pack_vertices(xyz, rgb), np.array([0, 1, 2], np.int64)


@pytest.mark.parametrize("flags", [0, O.LITERAL])
def test_kat1(flags):
    v, i = tri([(0, 1), (1, 0), (-1, 0)])
    rc, c, d, st = O.render(v, i, I4, 512, 512, flags=flags)
    assert rc == O.OK
    cov = c[..., 3] > 0
    assert cov.sum() == 256 * 256 + 1 == st["span_pixels"]
    for y in range(256):
        xs = np.nonzero(cov[y])[0]
        assert xs[0] == 256 - y and xs[-1] == 256 + y and xs.size == 2 * y + 1
    assert np.nonzero(cov[256])[0].tolist() == [0]
    assert c[0, 256].tolist() == [0, 0, 255, 255]
    assert c[128, 256].tolist() == [63, 63, 127, 255]
    assert c[256, 0].tolist() == [255, 0, 0, 255]
    assert c[100, 300].tolist() == [27, 71, 155, 255]
    assert (c[~cov] == 0).all()
    assert np.isposinf(d).all()


def test_kat2_flat_bottom_collapses_to_s2x():
    v, i = tri([(0, 0.5), (-0.5, -0.5), (0.5, -0.5)])
    rc, c, d, st = O.render(v, i, I4, 512, 512)
    cov = c[..., 3] > 0
    xs = np.nonzero(cov[383])[0]
    assert (xs[0], xs[-1]) == (129, 383)
    assert np.nonzero(cov[384])[0].tolist() == [384]
    assert not cov[385:].any()


def test_weights_match_kat():
    w = O.weights([(256, 0), (512, 256), (0, 256)], 300, 100)
    assert w.tolist() == [0.609375, 0.28125, 0.109375]


def test_interpolate_rules():
    pts = [(10, 0), (20, 10), (0, 30)]
    assert O.interpolate(pts, 0) == 10
    assert O.interpolate(pts, 5) == 15
    assert O.interpolate(pts, 10) == 20
    assert O.interpolate(pts, 29) == 20 + (-20 * 19) // 20          # exact
    assert O.interpolate(pts, 30) == 0 and O.interpolate(pts, 99) == 0   # t >= v2.y -> v2.x
    assert O.interpolate([(10, 0), (0, 3)], 1) == 10 + int(-10 * 1 / 3)  # truncation toward zero: -3, not -4
    assert O.interpolate([(10, 0), (0, 3)], 1) == 7
    assert O.interpolate([(7, 5), (99, 5)], 5) == 7                   # dy == 0 -> start.x


def test_pack_pixel():
    assert O.pack_pixel([1.0, 0.5, 0.0]).tolist() == [0, 127, 255, 255]
    assert O.pack_pixel([np.nan, -1.0, 2.0]).tolist() == [255, 0, 0, 255]
    assert O.pack_pixel([np.inf, -np.inf, 0.999]).tolist() == [254, 0, 255, 255]


def test_screen_mapping_not_rounded():
    s = O.transform_vertex(I4, [0.3003, -0.2, 0.25], 800, 600)
    assert s[0] == np.float32(np.float32(np.float32(0.3003) * np.float32(0.5) + np.float32(0.5)) * np.float32(800))
    assert s[1] == np.float32(np.float32(np.float32(-0.2) * np.float32(-0.5) + np.float32(0.5)) * np.float32(600))
    assert s[0] != np.round(s[0])


def test_all_equal_y_draws_min_max_of_s0_s2():
    # y identical after truncation: row span = [min(s0.x, s2.x), max(..)], s1.x is ignored
    v, i = tri([(-0.5, 0.0), (0.9, 0.0), (0.0, 0.0)])
    rc, c, d, st = O.render(v, i, I4, 100, 100)
    cov = c[..., 3] > 0
    assert cov.sum() == cov[50].sum() and np.nonzero(cov[50])[0].tolist() == list(range(25, 51))


def test_last_index_wins_and_depth_mode():
    xyz = np.array([[-1, -1, 0.5], [1, -1, 0.5], [0, 1, 0.5], [-1, -1, 0.25], [1, -1, 0.25], [0, 1, 0.25]], np.float32)
    rgb = np.array([[1, 0, 0]] * 3 + [[0, 1, 0]] * 3, np.float32)
    v = S.pack_vertices(xyz, rgb)
    a = np.array([0, 1, 2, 3, 4, 5], np.int64)
    b = np.array([3, 4, 5, 0, 1, 2], np.int64)
    _, ca, _, _ = O.render(v, a, I4, 64, 64)
    _, cb, _, _ = O.render(v, b, I4, 64, 64)
    assert ca[40, 32].tolist() == [0, 255, 0, 255] and cb[40, 32].tolist() == [0, 0, 255, 255]
    _, da, za, _ = O.render(v, a, I4, 64, 64, flags=O.DEPTH_LESS)
    _, db, zb, _ = O.render(v, b, I4, 64, 64, flags=O.DEPTH_LESS)
    assert np.array_equal(da, db) and np.array_equal(za, zb)
    assert da[40, 32].tolist() == [0, 255, 0, 255] and za[40, 32] == np.float32(0.25)


@pytest.mark.parametrize("seed", range(6))
def test_literal_equals_hoisted_random(seed):
    rng = np.random.default_rng(seed)
    n = 40
    xyz = rng.uniform(-1.3, 1.3, size=(3 * n, 3)).astype(np.float32)
    rgb = rng.uniform(-0.2, 1.2, size=(3 * n, 3)).astype(np.float32)
    v = S.pack_vertices(xyz, rgb)
    i = np.arange(3 * n, dtype=np.int64)
    for flags in (0, O.DEPTH_LESS):
        r1 = O.render(v, i, I4, 97, 61, flags=flags)
        r2 = O.render(v, i, I4, 97, 61, flags=flags | O.LITERAL)
        assert r1[0] == r2[0] == 0
        assert np.array_equal(r1[1], r2[1])
        assert np.array_equal(r1[2], r2[2])
        assert r1[3]["span_pixels"] == r2[3]["span_pixels"]


def test_vertices_and_line_primitives():
    v, i = tri([(0, 0), (0.5, 0.5), (5.0, 0.0)], colors=[[1, 0, 0], [0, 1, 0], [0, 0, 1]])
    rc, c, d, st = O.render(v, i, I4, 64, 64, primitive_type=O.VERTICES)
    assert rc == 0 and (c[..., 3] > 0).sum() == 2 and st["rejected_pixels"] == 1
    assert c[32, 32].tolist() == [0, 0, 255, 255] and c[16, 48].tolist() == [0, 255, 0, 255]
    rc, c, d, st = O.render(v, i, I4, 64, 64, primitive_type=O.LINE)
    assert rc == 0 and not c.any() and np.isposinf(d).all()


def test_error_conventions():
    v, i = tri([(0, 1), (1, 0), (-1, 0)])
    assert O.render(v, np.array([0, 1, 3]), I4, 8, 8)[0] == O.ERR_INDEX
    assert O.render(v, np.array([0, 1, 2, 7]), I4, 8, 8)[0] == O.OK        # trailing partial primitive ignored
    m = I4.copy()
    m[3, 3] = 0.0                                                            # w = 0 -> inf/NaN screen coords
    rc, c, d, st = O.render(v, i, m, 8, 8)
    assert rc == O.ERR_DOMAIN and st["domain_skipped"] == 1 and not c.any()


def test_mesh_generators_match_survey_sizes():
    v, i = S.cube()
    assert v.shape == (24, 8) and i.size == 36
    v, i = S.icosphere(3)
    assert i.size // 3 == 20 * 4 ** 3 and v.shape[0] == 10 * 4 ** 3 + 2
    v, i = S.cube_sphere(17)
    assert i.size // 3 == 12 * 17 * 17 and v.shape[0] == 6 * 17 * 17 + 2
    assert S.algorithmic_bytes(1_000_000, 3_000_000, 1920, 1080) == 100_588_800


def test_cube_scene_config1():
    """BASELINE.json configs[0]: rotating cube at 800x600 through the CPU path."""
    v, i = S.cube()
    prev = None
    for k in (0, 7, 30):
        rc, c, d, st = O.render(v, i, S.app_transform(k / 60), 800, 600)
        assert rc == 0 and st["primitives"] == 12 and (c[..., 3] > 0).sum() > 20000
        assert prev is None or not np.array_equal(prev, c)
        prev = c


def test_image_writers_roundtrip(tmp_path):
    import struct
    import zlib
    rng = np.random.default_rng(1)
    img = rng.integers(0, 256, size=(7, 5, 4), dtype=np.uint8)
    S.write_ppm(str(tmp_path / "a.ppm"), img)
    raw = (tmp_path / "a.ppm").read_bytes()
    assert raw.startswith(b"P6\n5 7\n255\n") and raw[11:] == img[..., 2::-1].tobytes()
    S.write_png(str(tmp_path / "a.png"), img)
    png = (tmp_path / "a.png").read_bytes()
    assert png[:8] == b"\x89PNG\r\n\x1a\n" and struct.unpack(">II", png[16:24]) == (5, 7)
    pos, idat = 8, b""
    while pos < len(png):
        n, tag = struct.unpack(">I", png[pos:pos + 4])[0], png[pos + 4:pos + 8]
        assert zlib.crc32(png[pos + 4:pos + 8 + n]) & 0xFFFFFFFF == struct.unpack(">I", png[pos + 8 + n:pos + 12 + n])[0]
        if tag == b"IDAT":
            idat += png[pos + 8:pos + 8 + n]
        pos += 12 + n
    rows = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(7, 1 + 5 * 4)
    assert (rows[:, 0] == 0).all() and np.array_equal(rows[:, 1:].reshape(7, 5, 4), img[..., [2, 1, 0, 3]])


def test_span_rule_is_invariant_under_row_shifts_but_weights_are_not():
    """The claim behind sr_set_band (screen-band split): interpolate (R:467-494) uses only differences of y, so moving
    the truncated corners AND the row by the same integer leaves every span end unchanged (both restatements: the
    C++ oracle and the pure-Python reading); ws(xy:) (R:88-100) is float arithmetic on Float(Int) + 0.5, which is
    NOT shift invariant for far off-screen corners, which is why the kernels shade in frame rows."""
    from oracle import literal_py as L
    rng = np.random.default_rng(21)
    moved = 0
    for _ in range(400):
        big = rng.random() < 0.3
        lim = (1 << 29) if big else 3000
        pts = rng.integers(-lim, lim, size=(3, 2))
        pts = pts[np.argsort(pts[:, 1], kind="stable")]              # sorted by y like R:271 leaves them
        shift = int(rng.integers(1, 2200))
        shifted = pts - np.array([0, shift])
        for vals in (pts, pts[[0, 2]]):                              # left = [s0, s1, s2], right = [s0, s2] (R:272-273)
            sv = vals - np.array([0, shift])
            for t in rng.integers(int(pts[0, 1]), int(pts[2, 1]) + 1, size=6):
                want = O.interpolate(vals, int(t))
                assert want == O.interpolate(sv, int(t) - shift)
                assert want == L.interpolate([tuple(map(int, p)) for p in vals], int(t))
                assert want == L.interpolate([tuple(map(int, p)) for p in sv], int(t) - shift)
        x, y = int(rng.integers(0, 1920)), int(rng.integers(shift, shift + 1080))
        w0 = O.weights(pts, x, y)
        w1 = O.weights(shifted, x, y - shift)
        if big and not np.array_equal(w0.view(np.uint32), w1.view(np.uint32)):
            moved += 1
        if not big:                                                  # on-screen magnitudes: every Float(Int) + 0.5 is exact
            assert np.array_equal(w0.view(np.uint32), w1.view(np.uint32))
    assert moved > 0
