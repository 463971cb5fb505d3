"""CPU check of the arithmetic claim behind csrc/sr_math.cuh::lerp_rcp (IEEE double on the host behaves like the
device's __drcp_rn / __dmul_rn):  for 0 <= n < 2^31, 1 <= d < 2^31,
    floor(n / d) == trunc(fl(double(n) * R)),   R = fl(fl(1/d) * (1 + 2^-40)).
This replaces the integer division of Renderer.swift:492 on the GPU, so it has to hold for every input."""
import numpy as np


def quot(n, d):
    R = (1.0 / d.astype(np.float64)) * (1.0 + 2.0 ** -40)
    return np.trunc(n.astype(np.float64) * R).astype(np.int64)


def check(n, d):
    n = np.asarray(n, np.int64)
    d = np.asarray(d, np.int64)
    ok = (n >= 0) & (n < 2 ** 31) & (d >= 1) & (d < 2 ** 31)
    n, d = n[ok], d[ok]
    assert np.array_equal(quot(n, d), n // d)


def test_exhaustive_small():
    d = np.arange(1, 2049, dtype=np.int64)
    n = np.arange(0, 1 << 15, dtype=np.int64)
    check(np.repeat(n, d.size), np.tile(d, n.size))


def test_random_and_adversarial():
    rng = np.random.default_rng(7)
    for _ in range(20):
        d = rng.integers(1, 2 ** 31, size=400_000)
        k = rng.integers(0, 2 ** 31, size=400_000) // d
        for off in (-1, 0, 1):                       # exact multiples and their neighbours
            check(k * d + off, d)
        check(rng.integers(0, 2 ** 31, size=400_000), d)
        small = rng.integers(1, 5000, size=400_000)    # screen-sized divisors, largest numerators
        check(2 ** 31 - 1 - rng.integers(0, 10_000, size=400_000), small)
        kk = (2 ** 31 - 1) // small
        check(kk * small, small)
        check(kk * small - 1, small)


def test_rect_touched_never_drops_a_covered_tile():
    """csrc/sr_math.cuh::rect_touched (bbox-tile pruning in k_setup) restated with Python ints: for random integer
    triangles, including flat tops/bottoms, slivers and repeated corners, a tile that holds a covered pixel of the
    span rule (Renderer.swift:275-283, 467-494) is never rejected; and the test does prune something."""
    from oracle import literal_py as P
    rng = np.random.default_rng(11)
    T, W, H = 16, 200, 160
    kept = pruned = 0
    for trial in range(3000):
        k = trial % 6
        pts = rng.integers(-40, 240, size=(3, 2))
        if k == 1:
            pts[1, 1] = pts[2, 1]                       # two corners on one row
        elif k == 2:
            pts[0, 1] = pts[1, 1] = pts[2, 1]           # all on one row
        elif k == 3:
            pts[1] = pts[0]                             # repeated corner
        elif k == 4:
            pts[:, 0] = pts[0, 0] + rng.integers(-2, 3, size=3)   # sliver
        order = sorted(range(3), key=lambda q: pts[q][1])       # stable, like R:271
        s = [(int(pts[q][0]), int(pts[q][1])) for q in order]
        y0, y1 = max(s[0][1], 0), min(s[2][1], H - 1)
        xmin, xmax = max(min(p[0] for p in s), 0), min(max(p[0] for p in s), W - 1)
        if y0 > y1 or xmin > xmax:
            continue
        spans = {}
        for y in range(y0, y1 + 1):
            a, b = P.interpolate(s, y), P.interpolate([s[0], s[2]], y)
            spans[y] = (min(a, b), max(a, b))
        for ty in range(y0 // T, y1 // T + 1):
            for tx in range(xmin // T, xmax // T + 1):
                ya, yb = max(ty * T, s[0][1], 0), min(ty * T + T - 1, s[2][1], H - 1)
                x_lo, x_hi = tx * T, min(tx * T + T - 1, W - 1)
                covered = any(spans[y][0] <= x_hi and spans[y][1] >= x_lo for y in range(ya, yb + 1))
                mn, mx = min(spans[ya][0], spans[yb][0]), max(spans[ya][1], spans[yb][1])
                if ya < s[1][1] <= yb:
                    mn, mx = min(mn, s[1][0]), max(mx, s[1][0])
                keep = mn <= x_hi and mx >= x_lo
                assert keep or not covered, (s, tx, ty)
                kept += keep
                pruned += not keep
    assert pruned > 0.2 * (kept + pruned)


def test_rank_ordered_round_of_the_tile_kernel():
    """csrc/sr_kernels.cuh::k_tile, depth off, restated with Python ints: the 16 triangles of a round are filed by
    descending primitive index (rank = how many of the round have a higher index) in column ((rank & 7) << 1) |
    (rank >> 3) of the 16 x 16-bit row masks; a pixel gathers bit q from the low half of word q and bit 16 + q from
    its high half, and takes the LOWEST set bit as the winner.  For random rounds (any length, any coverage) that
    must be the highest covering index — the outcome of the sequential loop Renderer.swift:221-229 with the
    unconditional store of R:268 — and the early-out threshold must come from the round's lowest index."""
    rng = np.random.default_rng(11)
    for _ in range(3000):
        n = int(rng.integers(1, 17))
        prims = rng.choice(1 << 20, size=n, replace=False)              # a tile list holds a triangle once
        cover = rng.random(n) < rng.random()                            # which triangles cover the pixel
        col_mask = np.zeros(16, dtype=np.int64)                         # one pixel: bit 0 of each column's row mask
        s_prim = np.full(16, -1, dtype=np.int64)
        for j in range(16):
            rank = int((prims > prims[j]).sum()) if j < n else j
            col = ((rank & 7) << 1) | (rank >> 3)
            assert col_mask[col] == 0 and (j >= n or s_prim[col] == -1)  # ranks are a permutation: no column is shared
            if j < n:
                s_prim[col] = prims[j]
                col_mask[col] = int(cover[j])
        words = [(int(col_mask[2 * q]) | (int(col_mask[2 * q + 1]) << 16)) for q in range(8)]
        bits = 0
        for q in range(8):
            bits |= (words[q] & 0x00010001) << q
        if cover.any():
            b = (bits & -bits).bit_length() - 1
            assert s_prim[((b & 15) << 1) | (b >> 4)] == prims[cover].max()
        else:
            assert bits == 0
        assert s_prim[(((n - 1) & 7) << 1) | ((n - 1) >> 3)] == prims.min()


def test_row_has_open_bit_logic():
    """csrc/sr_kernels.cuh::row_has_open (retired-tile bitmap, one bit per tile, 32 tiles per word) restated with
    Python ints against the definition: some tile column in [txa, txb] has its 'done' bit clear."""
    rng = np.random.default_rng(5)
    M = 0xFFFFFFFF
    for _ in range(20000):
        ncols = int(rng.integers(1, 200))
        wpr = (ncols + 31) // 32
        density = rng.choice([0.0, 0.5, 0.97, 1.0])
        done = rng.random(ncols) < density
        words = [0] * wpr
        for tx in range(ncols):
            if done[tx]:
                words[tx >> 5] |= 1 << (tx & 31)
        txa = int(rng.integers(0, ncols))
        txb = int(rng.integers(txa, ncols))
        wa, wb = txa >> 5, txb >> 5
        first, last = (M << (txa & 31)) & M, M >> (31 - (txb & 31))
        if wa == wb:
            got = ((~words[wa]) & M & first & last) != 0
        else:
            got = ((~words[wa]) & M & first) != 0 or any(((~words[w]) & M) != 0 for w in range(wa + 1, wb)) \
                or ((~words[wb]) & M & last) != 0
        assert got == (not done[txa:txb + 1].all())


def test_float32_quotient_with_one_remainder_step_is_exact():
    """csrc/sr_math.cuh lerp_f32 (-DSR_DIV32 experiment): for 0 <= a < 2^22, 1 <= d < 2^24 the float estimate
    trunc(fl(float(a) * r)) is within one of floor(a / d) for EVERY r within rcp.approx's relative error 2^-23 of 1/d, and
    one remainder step repairs it.  The host cannot reproduce the GPU's rcp.approx bit for bit, so both ends of its error
    interval (and the exact reciprocal) are checked: any value in between gives an estimate between these."""
    rng = np.random.default_rng(7)

    def check(a, d):
        a = np.asarray(a, np.int64)
        d = np.asarray(d, np.int64)
        want = a // d
        af, df = a.astype(np.float32), d.astype(np.float32)
        assert np.array_equal(af.astype(np.int64), a) and np.array_equal(df.astype(np.int64), d)      # both exact in float32
        exact = 1.0 / df.astype(np.float64)
        for scale in (1.0 - 2.0 ** -23, 1.0, 1.0 + 2.0 ** -23):
            r = (exact * scale).astype(np.float32)
            q0 = np.trunc((af * r).astype(np.float32)).astype(np.int64)        # one float32 product, truncated
            assert np.abs(q0 - want).max() <= 1
            rem = a - q0 * d
            assert np.abs(rem).max() < 2 ** 25                                  # the 32-bit remainder cannot overflow
            q = q0 + (rem >= d) - (rem < 0)
            assert np.array_equal(q, want)

    # exhaustive over small operands
    aa, dd = np.meshgrid(np.arange(0, 600), np.arange(1, 300))
    check(aa.ravel(), dd.ravel())
    # random over the whole admitted range, and the adversarial cases: exact multiples and one below
    d = rng.integers(1, 1 << 24, 400_000)
    a = rng.integers(0, 1 << 22, 400_000)
    check(a, d)
    d = rng.integers(1, 1 << 11, 400_000)
    k = rng.integers(0, 1 << 11, 400_000)
    for delta in (0, -1, 1):
        a = np.clip(k * d + delta, 0, (1 << 22) - 1)
        check(a, d)
    check(np.full(4096, (1 << 22) - 1), np.arange(1, 4097))                     # largest numerator


def test_late_box_word_equals_the_full_setups_tile_range():
    """csrc/sr_kernels.cuh::tile_box_word (k_late_boxes) restated with numpy: the clipped tile box taken from the min / max of
    the TRUNCATED corners equals the one the full setup derives after its stable 3-sort on the float rows (truncation toward
    zero is monotone, so the sorted corners' rows are the min / max of the truncated rows); the word's second byte is 0xFF
    only for 'nothing on screen'; bands (y_org) shift rows before clipping."""
    rng = np.random.default_rng(0xB0C5)
    for W, H, y_org in [(1920, 1080, 0), (4080, 4080, 0), (333, 251, 0), (1920, 135, 405), (17, 9, 3)]:
        n = 20000
        scale = rng.choice([0.3, 1.0, 3.0, 50.0], size=(n, 1, 1))
        c = rng.uniform(-0.4, 1.4, size=(n, 1, 2)) * np.array([W, H + y_org])
        xy = (c + rng.normal(size=(n, 3, 2)) * scale * 40.0).astype(np.float32)
        xy[rng.random(n) < 0.05] *= np.float32(-1.0)                       # negative coordinates: truncation goes UP
        ix = np.trunc(xy[..., 0]).astype(np.int64)
        iy = np.trunc(xy[..., 1]).astype(np.int64)
        # the box pre-pass: min / max of the truncated corners, rows moved to the band, clipped
        ya = np.maximum(iy.min(1) - y_org, 0); yb = np.minimum(iy.max(1) - y_org, H - 1)
        xa = np.maximum(ix.min(1), 0); xb = np.minimum(ix.max(1), W - 1)
        on = (ya <= yb) & (xa <= xb)
        word = (xa >> 4) | ((xb >> 4) << 8) | ((ya >> 4) << 16) | ((yb >> 4) << 24)
        word = np.where(on, word, 0xFFFFFFFE)
        # the full setup: stable insertion sort with strict < on the FLOAT rows, then truncate and clip (R:271-273)
        order = np.argsort(xy[..., 1], axis=1, kind="stable")
        sy = np.take_along_axis(iy, order, 1) - y_org
        y0 = np.maximum(sy[:, 0], 0); y1 = np.minimum(sy[:, 2], H - 1)
        assert np.array_equal(on, (y0 <= y1) & (xa <= xb))
        full = (xa >> 4) | ((xb >> 4) << 8) | ((y0 >> 4) << 16) | ((y1 >> 4) << 24)
        assert np.array_equal(word[on], full[on])
        assert not (((word[on] >> 8) & 0xFF) == 0xFF).any() and (((word[~on] >> 8) & 0xFF) == 0xFF).all()
