// sr_math.cuh — the arithmetic of Renderer.swift, as device functions.
//
// Every float operation that decides a coverage bit, a depth bit or a colour byte is written with the
// round-to-nearest intrinsics (__fmul_rn, __fadd_rn, __fsub_rn, __fdiv_rn), which nvcc never contracts into FMA,
// in the reference's operation order.  The translation unit is also built with -fmad=false.
// "R:" = /root/reference/renderer/Renderer.swift.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace sr {

struct Mat4 { float m[16]; };  // column-major, like matrix_float4x4 (R:199)

// simd_muladd(a, b, c), the accumulation step of Apple's simd_mul (R:160, R:91): a*b + c rounded twice by default (x86
// without FMA; "FMA contraction disabled"), one fused multiply-add with -DSR_SIMD_FMA (arm64, the platform the
// reference ships on).  libsr_cuda_fma.so is the second build; oracle/ has the same switch.
__device__ __forceinline__ float simd_muladd(float a, float b, float c) {
#ifdef SR_SIMD_FMA
    return __fmaf_rn(a, b, c);
#else
    return __fadd_rn(__fmul_rn(a, b), c);
#endif
}

// Vertex.apply (R:159-163) + convertedToScreen (R:165-172).  Returns (sx, sy, z_ndc).
// matrix * float4 is column accumulation with separately rounded * and + (the definition shared with oracle/).
// WITH_Z = false leaves z_ndc out (returned as 0): with the depth test off nothing reads it.
template <bool WITH_Z = true>
__device__ __forceinline__ float3 transform_to_screen(const Mat4& M, float x, float y, float z, float fw, float fh) {
    float v[4];
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        if (!WITH_Z && r == 2) { v[r] = 0.0f; continue; }
        float acc = __fmul_rn(M.m[r], x);
        acc = simd_muladd(M.m[4 + r], y, acc);
        acc = simd_muladd(M.m[8 + r], z, acc);
        acc = simd_muladd(M.m[12 + r], 1.0f, acc);
        v[r] = acc;
    }
    float nx = __fdiv_rn(v[0], v[3]);   // R:161: three true divisions, no reciprocal
    float ny = __fdiv_rn(v[1], v[3]);
    float nz = WITH_Z ? __fdiv_rn(v[2], v[3]) : 0.0f;
    float ux = __fadd_rn(__fmul_rn(nx, 0.5f), 0.5f);    // R:166
    float uy = __fadd_rn(__fmul_rn(ny, -0.5f), 0.5f);
    return make_float3(__fmul_rn(ux, fw), __fmul_rn(uy, fh), nz);   // R:168 (no rounding of the product)
}

// valid domain: finite and |s| < 2^30 (NaN fails).  Outside it the reference traps (R:251, R:271, R:298).
__device__ __forceinline__ bool in_domain(float s) { return fabsf(s) < 1073741824.0f; }

// ---- integer span rule (R:467-494) ----------------------------------------------------------------------------
// start.x + diff*nt/dy, 64-bit product, division truncating toward zero.  dy != 0.
__device__ __forceinline__ int lerp_trunc(int sx, int sy, int ex, int ey, int t) {
    long long diff = (long long)ex - (long long)sx;
    long long nt = (long long)t - (long long)sy;
    long long dy = (long long)ey - (long long)sy;
    long long num = diff * nt;
    long long q;
    if (num == (long long)(int)num && dy == (long long)(int)dy && !((int)num == INT32_MIN))
        q = (long long)((int)num / (int)dy);      // 32-bit fast path (always taken for on-screen geometry)
    else
        q = num / dy;
    return (int)((long long)sx + q);
}

struct SortedTri { int x0, y0, x1, y1, x2, y2; };   // s0, s1, s2 of R:271 (y0 <= y1 <= y2)

// interpolate(values: leftVertices, t:)  — the 3-point list [s0,s1,s2] (R:272, R:469-475)
// Rows whose quotient is known without dividing take a shortcut: nt == 0 gives start.x, t >= y2 gives s2.x.
__device__ __forceinline__ int span_left(const SortedTri& s, int t) {
    if (t >= s.y2) return s.x2;                           // R:470-471 + R:478-480: last row collapses to s2.x
    if (t >= s.y1) return t == s.y1 ? s.x1 : lerp_trunc(s.x1, s.y1, s.x2, s.y2, t);
    return t == s.y0 ? s.x0 : lerp_trunc(s.x0, s.y0, s.x1, s.y1, t);
}
// interpolate(values: rightVertices, t:) — the 2-point list [s0,s2] (R:273)
__device__ __forceinline__ int span_right(const SortedTri& s, int t) {
    if (s.y2 == s.y0 || t == s.y0) return s.x0;           // R:486-488; nt == 0
    if (t == s.y2) return s.x2;                           // diff*dy/dy is exact
    return lerp_trunc(s.x0, s.y0, s.x2, s.y2, t);
}
// One row of R:275-283: returns inclusive [l, r] after the swap of R:278-280.
__device__ __forceinline__ void span_row(const SortedTri& s, int t, int& l, int& r) {
    int a = span_left(s, t), b = span_right(s, t);
    l = min(a, b);
    r = max(a, b);
}

// ---- the same rule with per-edge reciprocals ---------------------------------------------------------------------
// For integers 0 <= n < 2^31 and 1 <= d < 2^31:  floor(n / d) == trunc(fl(double(n) * R)),  R = fl(fl(1/d) * (1 + 2^-40)).
// R = (1/d)(1 + e) with e in [2^-40 - 2^-51, 2^-40 + 2^-51]; the rounded product is (n/d)(1 + e)(1 + g), |g| <= 2^-53,
// which is >= n/d (so an exact quotient never drops below its integer) and exceeds n/d by less than
// n * 2^-39 / d < 1/d, the distance from n/d to the next integer boundary (so the floor never moves up).
// R is computed once per edge in triangle setup; a row then costs one conversion, one DMUL and one truncation
// instead of an integer division.  Larger numerators (far off-screen corners) fall back to the 64-bit division,
// so the result is always the reference's quotient.  tests/test_host_math.py checks the claim on the CPU.
__device__ __forceinline__ double edge_rcp(int dy) {
    return dy > 0 ? __dmul_rn(__drcp_rn((double)dy), 1.0 + 0x1p-40) : 0.0;
}
__device__ __forceinline__ int lerp_rcp(int sx, int sy, int ex, int ey, int t, double R) {
    const long long num = (long long)(ex - sx) * (long long)(t - sy);       // |ex - sx|, |t - sy| < 2^31 inside the domain
    const unsigned long long a = (unsigned long long)(num < 0 ? -num : num);
    unsigned long long q;
    if (a < (1ull << 31)) q = (unsigned long long)__double2uint_rz(__dmul_rn((double)(unsigned)a, R));
    else q = a / (unsigned long long)((long long)ey - (long long)sy);
    return (int)((long long)sx + (num < 0 ? -(long long)q : (long long)q));
}
// ---- the same quotient in 32-bit float arithmetic (-DSR_DIV32 experiment) ---------------------
// For integers 0 <= a < 2^22 and 1 <= d < 2^24:  q0 = trunc(fl(float(a) * r)) with r = rcp.approx(float(d)) is within
// one of floor(a / d), and one remainder step makes it exact.  float(a), float(d) are exact; rcp.approx has a relative
// error <= 2^-23 and the product rounds once more (2^-24), so |q0' - a/d| <= (a/d) * 1.5 * 2^-23 < 2^22 * 1.5 * 2^-23 =
// 0.75 before truncation: the truncated value is floor(a/d) - 1, floor(a/d) or floor(a/d) + 1.  With rem = a - q0 * d
// (|rem| < 2 d < 2^25, no overflow) the quotient is q0 + (rem >= d) - (rem < 0).  Anything larger takes the 64-bit
// division, so the result is always the reference's (R:492).  tests/test_host_math.py checks the claim on the CPU for
// every reciprocal within the stated error.
__device__ __forceinline__ float rcp_approx(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ int lerp_f32(int sx, int sy, int ex, int ey, int t) {
    const long long num = (long long)(ex - sx) * (long long)(t - sy);
    const unsigned long long a = (unsigned long long)(num < 0 ? -num : num);
    const long long dy = (long long)ey - (long long)sy;
    unsigned long long q;
    if (a < (1ull << 22) && dy < (1ll << 24)) {
        const unsigned au = (unsigned)a, d = (unsigned)dy;
        unsigned q0 = __float2uint_rz(__fmul_rn(__uint2float_rn(au), rcp_approx(__uint2float_rn(d))));
        const int rem = (int)au - (int)(q0 * d);
        q0 += (rem >= (int)d) ? 1u : 0u;
        q0 -= (rem < 0) ? 1u : 0u;
        q = q0;
    } else {
        q = a / (unsigned long long)dy;
    }
    return (int)((long long)sx + (num < 0 ? -(long long)q : (long long)q));
}
__device__ __forceinline__ void span_row_f32(const SortedTri& s, int t, int& l, int& r) {
    int a, b;
    if (t >= s.y2) a = s.x2;
    else if (t >= s.y1) a = lerp_f32(s.x1, s.y1, s.x2, s.y2, t);
    else a = lerp_f32(s.x0, s.y0, s.x1, s.y1, t);
    b = (s.y2 == s.y0) ? s.x0 : lerp_f32(s.x0, s.y0, s.x2, s.y2, t);
    l = min(a, b);
    r = max(a, b);
}

struct EdgeRcp { double r01, r12, r02; };
__device__ __forceinline__ void span_row_rcp(const SortedTri& s, const EdgeRcp& m, int t, int& l, int& r) {
    int a, b;
    if (t >= s.y2) a = s.x2;
    else if (t >= s.y1) a = lerp_rcp(s.x1, s.y1, s.x2, s.y2, t, m.r12);
    else a = lerp_rcp(s.x0, s.y0, s.x1, s.y1, t, m.r01);
    b = (s.y2 == s.y0) ? s.x0 : lerp_rcp(s.x0, s.y0, s.x2, s.y2, t, m.r02);
    l = min(a, b);
    r = max(a, b);
}

// Can triangle `s` cover a pixel of the rectangle columns [x_lo, x_hi] x rows [ya, yb] (ya..yb inside the triangle's
// own row range)?  Both span ends are monotone in the row on each edge piece (x + trunc(diff*nt/dy) is monotone in nt,
// and the pieces join continuously, except that a flat-bottomed triangle (y1 == y2) jumps from near s1.x to s2.x on
// its last row, R:470-471), so over the rows ya..yb the smallest left end / largest right end are reached at ya, yb,
// or are bounded by s1.x when the middle corner's row lies in (ya, yb].  The test "min end <= x_hi and max end >= x_lo" never rejects
// a rectangle that holds a covered pixel; it may keep one that holds none (the two extremes can sit on different rows).
__device__ __forceinline__ void row_band_extent(const SortedTri& s, const EdgeRcp& m, int ya, int yb, int& mn, int& mx) {
    int l0, r0, l1, r1;
#ifdef SR_DIV32_SETUP
    span_row_f32(s, ya, l0, r0);         // experiment: the same quotient as k_tile's, no per-edge FP64 reciprocals
    span_row_f32(s, yb, l1, r1);
#else
    span_row_rcp(s, m, ya, l0, r0);
    span_row_rcp(s, m, yb, l1, r1);
#endif
    mn = min(l0, l1);
    mx = max(r0, r1);
    if (s.y1 > ya && s.y1 <= yb) { mn = min(mn, s.x1); mx = max(mx, s.x1); }  // left chain at / towards its corner row
}
__device__ __forceinline__ bool rect_touched(const SortedTri& s, const EdgeRcp& m, int x_lo, int x_hi, int ya, int yb) {
    int mn, mx;
    row_band_extent(s, m, ya, yb, mn, mx);
    return mn <= x_hi && mx >= x_lo;
}

// Stable insertion sort of 3 keys with strict < (Swift's sort on 3 elements, R:271).  Returns the permutation
// as three 2-bit fields: perm = i0 | i1<<2 | i2<<4.
__device__ __forceinline__ unsigned sort3_perm(float ya, float yb, float yc) {
    int i0 = 0, i1 = 1, i2 = 2;
    float k0 = ya, k1 = yb, k2 = yc;
    if (k1 < k0) { float t = k0; k0 = k1; k1 = t; int u = i0; i0 = i1; i1 = u; }
    if (k2 < k1) {
        { float t = k1; k1 = k2; k2 = t; int u = i1; i1 = i2; i2 = u; }
        if (k1 < k0) { float t = k0; k0 = k1; k1 = t; int u = i0; i0 = i1; i1 = u; }
    }
    return (unsigned)i0 | ((unsigned)i1 << 2) | ((unsigned)i2 << 4);
}

// ---- barycentric weights (R:88-100) ---------------------------------------------------------------------------
struct BaryT { float c0x, c0y, c1x, c1y, cfx, cfy; };   // columns of T() = inverse, and cf

// T() with the 2x2 inverse defined as adjugate / det, det and every quotient rounded separately, no singular
// early-out (shared definition with oracle/; Apple's __invert_f2 is closed source).
__device__ __forceinline__ BaryT bary_setup(int ax, int ay, int bx, int by, int cx, int cy) {
    float afx = __fadd_rn(__int2float_rn(ax), 0.5f), afy = __fadd_rn(__int2float_rn(ay), 0.5f);
    float bfx = __fadd_rn(__int2float_rn(bx), 0.5f), bfy = __fadd_rn(__int2float_rn(by), 0.5f);
    float cfx = __fadd_rn(__int2float_rn(cx), 0.5f), cfy = __fadd_rn(__int2float_rn(cy), 0.5f);
    float m00 = __fsub_rn(afx, cfx), m10 = __fsub_rn(afy, cfy);   // column 0
    float m01 = __fsub_rn(bfx, cfx), m11 = __fsub_rn(bfy, cfy);   // column 1
    float det = __fsub_rn(__fmul_rn(m00, m11), __fmul_rn(m01, m10));
    BaryT t;
    t.c0x = __fdiv_rn(m11, det);
    t.c0y = __fdiv_rn(-m10, det);
    t.c1x = __fdiv_rn(-m01, det);
    t.c1y = __fdiv_rn(m00, det);
    t.cfx = cfx;
    t.cfy = cfy;
    return t;
}

// ws(xy:) at the centre of pixel (x, y)  (R:252, R:88-93)
__device__ __forceinline__ float3 bary_weights(const BaryT& t, int x, int y) {
    float px = __fadd_rn(__int2float_rn(x), 0.5f), py = __fadd_rn(__int2float_rn(y), 0.5f);
    float dx = __fsub_rn(px, t.cfx), dy = __fsub_rn(py, t.cfy);
    float wa = simd_muladd(t.c1x, dy, __fmul_rn(t.c0x, dx));     // simd_mul(float2x2, float2), R:91
    float wb = simd_muladd(t.c1y, dy, __fmul_rn(t.c0y, dx));
    return make_float3(wa, wb, __fsub_rn(__fsub_rn(1.0f, wa), wb));
}

// a*w.x + b*w.y + c*w.z, left to right, unfused (R:257, R:266)
__device__ __forceinline__ float blend3(float a, float b, float c, float3 w) {
    return __fadd_rn(__fadd_rn(__fmul_rn(a, w.x), __fmul_rn(b, w.y)), __fmul_rn(c, w.z));
}

// Pixel(float3:) (R:117-128): clamp = fmin(fmax(v,0),1) (NaN -> 0), * 255, truncate; bytes b,g,r,a.
__device__ __forceinline__ unsigned pack_bgra(float r, float g, float b) {
    unsigned ub = (unsigned)__float2int_rz(__fmul_rn(fminf(fmaxf(b, 0.0f), 1.0f), 255.0f));
    unsigned ug = (unsigned)__float2int_rz(__fmul_rn(fminf(fmaxf(g, 0.0f), 1.0f), 255.0f));
    unsigned ur = (unsigned)__float2int_rz(__fmul_rn(fminf(fmaxf(r, 0.0f), 1.0f), 255.0f));
    return ub | (ug << 8) | (ur << 16) | 0xFF000000u;
}

// ---- depth keys (SR_FLAG_DEPTH_LESS; R:257-261) ----------------------------------------------------------------
// Sequential rule: a pixel keeps the smallest depth; among equal depths the earliest primitive (strict <).
// key = monotone(depth) << 32 | prim, resolved with atomicMin.  NaN and +inf never pass `depth < +inf`.
#define SR_KEY_EMPTY 0xFFFFFFFFFFFFFFFFull
__device__ __forceinline__ bool depth_can_pass(float d) { return d < __int_as_float(0x7f800000); }
__device__ __forceinline__ unsigned long long depth_key(float d, unsigned prim) {
    unsigned u = __float_as_uint(d == 0.0f ? 0.0f : d);   // -0 and +0 compare equal in the reference's test
    u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);
    return ((unsigned long long)u << 32) | prim;
}

// ---- conservative depth bounds (depth mode: hierarchical rejection) ----------------------------------------------
// ordered-uint image of a float (monotone: a < b  =>  ord(a) < ord(b); -0 sorts just below +0)
__device__ __forceinline__ unsigned depth_ord(float d) {
    const unsigned u = __float_as_uint(d);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float depth_unord(unsigned o) {
    return __uint_as_float((o & 0x80000000u) ? (o & 0x7fffffffu) : ~o);
}
// One corner sample for depth_lower_bound: the depth R:257 computes at pixel (x, y), and the magnitudes that scale
// its rounding error.
struct DepthSample { float d, a, b; };
__device__ __forceinline__ DepthSample depth_sample(const BaryT& t, float za, float zb, float zc, int x, int y) {
    const float px = __fadd_rn(__int2float_rn(x), 0.5f), py = __fadd_rn(__int2float_rn(y), 0.5f);
    const float dx = __fsub_rn(px, t.cfx), dy = __fsub_rn(py, t.cfy);
    DepthSample s;
    s.d = blend3(za, zb, zc, bary_weights(t, x, y));
    s.a = fabsf(t.c0x * dx) + fabsf(t.c1x * dy);
    s.b = fabsf(t.c0y * dx) + fabsf(t.c1y * dy);
    return s;
}
// Lower bound of every depth value R:257 can produce for this triangle at the pixels of a rectangle, given the
// samples at the rectangle's four corner pixels (dmin = smallest sample, amax / bmax = largest magnitudes).
// In real arithmetic the depth is affine in (x, y) (the per-triangle constants are fixed floats), so its minimum
// over the rectangle sits at a corner; the float evaluation differs from the real one by at most
// ~8u (|za| A + |zb| B + |zc| (1 + A + B)), u = 2^-24, A = |c0x dx| + |c1x dy|, B likewise (each product, sum and
// the two subtractions of wc round once; A and B are convex, so their maxima are at corners too).  The margin below
// uses 2^-17 (16x that) twice: once for the interior pixel, once for the corner sample.  A non-finite result means
// "no bound" and callers must not reject on it.
__device__ __forceinline__ float depth_lower_bound(float dmin, float amax, float bmax, float za, float zb, float zc) {
    const float e = (fabsf(za) * amax + fabsf(zb) * bmax + fabsf(zc) * (1.0f + amax + bmax)) * 0x1p-16f;
    return dmin - e - 1e-37f;
}

// Lower bound of the depth at EVERY pixel the span rule can cover for this triangle.  Each covered pixel centre lies
// on a row that crosses the real triangle (af, bf, cf) and within one pixel, horizontally, of that crossing (both
// span ends are truncations of the exact edge positions, or a vertex on the collapsed last row), so the depth there
// is at least min(za, zb, zc) - |dD/dx| - rounding, with dD/dx = za c0x + zb c0y - zc (c0x + c0y).
// amax / bmax: the DepthSample magnitudes maximised over a rectangle that holds all covered pixels (clipped bbox).
__device__ __forceinline__ float depth_triangle_bound(const BaryT& t, float za, float zb, float zc, float amax, float bmax) {
    const float gx = (za * t.c0x + zb * t.c0y) - zc * (t.c0x + t.c0y);
    const float slack = (fabsf(za * t.c0x) + fabsf(zb * t.c0y) + fabsf(zc * t.c0x) + fabsf(zc * t.c0y)) * 0x1p-16f;
    const float e = (fabsf(za) * amax + fabsf(zb) * bmax + fabsf(zc) * (1.0f + amax + bmax)) * 0x1p-16f;
    return fminf(za, fminf(zb, zc)) - fabsf(gx) - slack - e - 1e-37f;
}


// Rigorous lower bound of every depth value R:257 can produce for the triangle (original corner order a, b, c, FRAME
// coordinates) at a pixel the span rule covers inside the clipped bounding box [xmin, xmax] x [ymin, ymax] (frame
// rows): the better of the rectangle bound and the covered-pixel bound.  NaN = "no bound" (never reject on it).
__device__ __forceinline__ float triangle_depth_floor(int ax, int ay, int bx, int by, int cx, int cy, float za, float zb, float zc,
                                                      int xmin, int ymin, int xmax, int ymax) {
    const BaryT bt = bary_setup(ax, ay, bx, by, cx, cy);
    DepthSample m = depth_sample(bt, za, zb, zc, xmin, ymin);
#pragma unroll
    for (int q = 1; q < 4; ++q) {
        const DepthSample t = depth_sample(bt, za, zb, zc, (q & 1) ? xmax : xmin, (q & 2) ? ymax : ymin);
        m.d = fminf(m.d, t.d); m.a = fmaxf(m.a, t.a); m.b = fmaxf(m.b, t.b);
    }
    float dlo = fmaxf(depth_lower_bound(m.d, m.a, m.b, za, zb, zc), depth_triangle_bound(bt, za, zb, zc, m.a, m.b));
    if (!(fabsf(m.d) < __int_as_float(0x7f800000) && fabsf(m.a + m.b) < __int_as_float(0x7f800000))) dlo = __int_as_float(0x7fc00000);
    return dlo;
}

}  // namespace sr
