/*
 * sr_oracle.h — CPU restatement of zhvrnkov/software-renderer's Renderer.render(renderPass:)
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product path.  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it, and
 * only as the checker / timed CPU baseline.  libsr_cuda.so never links or calls it.
 *
 * PARITY UNPINNED: the reference is Swift (renderer/Renderer.swift) and there is no Swift
 * toolchain in this image, so the reference cannot be compiled or run here, and it ships no tests,
 * fixtures or golden vectors.  This restatement follows the reference line by line
 * (citations below, all into /root/reference/renderer/Renderer.swift = "R:") and is anchored on two
 * hand-derived known-answer vectors (KAT-1, KAT-2; tests/test_oracle_kat.py).  Two operations live in
 * Apple's closed `simd` library and are *defined* here (shared verbatim by the CUDA path):
 *   (1) matrix_float4x4 * simd_float4   (R:160): column accumulation, every * and + rounded
 *       separately (no FMA):  r = c0*x; r = r + c1*y; r = r + c2*z; r = r + c3*w.
 *   (2) matrix_float2x2.inverse         (R:99):  det = m00*m11 - m01*m10 (unfused); every adjugate
 *       entry is divided by det with a true IEEE division; no singular early-out (inf/NaN flow on).
 *   (3) simd_clamp(v,0,1)               (R:119-122): fminf(fmaxf(v,0),1)  (NaN -> 0).
 *
 * Valid input domain (the reference *traps* outside it: R:251, R:271, R:298, R:492):
 *   a primitive is drawn only if every corner's screen x,y is finite and |.| < 2^30; otherwise the
 *   primitive (for .vertices: the corner) is skipped and counted in stats.domain_skipped, and the
 *   call returns SR_ORACLE_ERR_DOMAIN after rendering everything else.
 */
#ifndef SR_ORACLE_H
#define SR_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Same layout as Swift's `Vertex` (R:154-157): two SIMD3<Float>, each padded to 16 bytes. */
typedef struct sr_oracle_vertex {
    float x, y, z, _pad0;
    float r, g, b, _pad1;
} sr_oracle_vertex;

/* PrimitiveType (R:174-189) */
enum { SR_ORACLE_TRIANGLE = 0, SR_ORACLE_LINE = 1, SR_ORACLE_VERTICES = 2 };

/* flags */
enum {
    SR_ORACLE_DEPTH_LESS = 1u << 0, /* R:257-261 uncommented (strict <, earliest index wins ties) */
    SR_ORACLE_LITERAL    = 1u << 8  /* per-pixel recomputation and unclipped loops exactly as written */
};

enum { SR_ORACLE_OK = 0, SR_ORACLE_ERR_ARG = -1, SR_ORACLE_ERR_INDEX = -2, SR_ORACLE_ERR_DOMAIN = -3 };

typedef struct sr_oracle_stats {
    uint64_t primitives;      /* primitives looped over (R:222) */
    uint64_t domain_skipped;  /* primitives (or .vertices corners) outside the valid domain */
    uint64_t span_pixels;     /* setPixel calls that passed the bounds guard (R:246-250), incl. overdraw */
    uint64_t rejected_pixels; /* setPixel calls rejected by the guard (LITERAL only; the reference prints) */
    uint64_t depth_passed;    /* DEPTH_LESS: pixels that passed the test */
} sr_oracle_stats;

/*
 * render(renderPass:)  R:204-230.
 *   transform: 16 floats, column-major (columns[c][r] = m[c*4+r]) like matrix_float4x4.
 *   indices:   Swift `[Int]`, 8 bytes each.  A trailing partial primitive is ignored (R:221).
 *   color:     width*height BGRA8 (R:44-49), packed y*width+x (App.swift:351-360).
 *   depth:     width*height float.
 */
int sr_oracle_render(const sr_oracle_vertex* vertices, size_t nverts,
                     const int64_t* indices, size_t nidx,
                     const float* transform, int primitive_type,
                     int width, int height,
                     uint8_t* color_bgra, float* depth,
                     uint32_t flags, sr_oracle_stats* stats);

/* Exposed pieces, for unit tests of the stage kernels. */
/* Vertex.apply + convertedToScreen (R:159-172): out = (sx, sy, z). */
void sr_oracle_transform_vertex(const float* transform, const float* xyz, int width, int height, float* out3);
/* interpolate (R:467-494) on 3-point (n=3) or 2-point (n=2) list of (x,y) int64 pairs. */
int64_t sr_oracle_interpolate(const int64_t* xy_pairs, int n, int64_t t);
/* Triangle.ws (R:88-100) for integer corners a,b,c and pixel (x,y): out = ws.xyz */
void sr_oracle_weights(const int64_t* abc_xy, int64_t x, int64_t y, float* out3);
/* Pixel(float3:) (R:117-128): rgb floats -> 4 bytes b,g,r,a */
void sr_oracle_pack_pixel(const float* rgb, uint8_t* out_bgra);

/*
 * Legacy 2-D immediate-mode API of the same file: draw(rect:) R:376-383, draw(circle:)
 * R:385-403, draw(line:) R:405-419, draw(triangle:with:in:) R:421-465.  Commands run in order on one BGRA image.
 *   RECT     p = {x, y, w, h}: rows y ..< y+w step sign(w), columns x ..< x+h step sign(h) (w and h are swapped in
 *            the reference, R:378-379), raw pointer stores of `color` (no bounds check).
 *   CIRCLE   p = {x, y, r}: dy, dx in -r ..< r; d = 1 if |p| < r-2 else clamp(smoothstep(r+2, r-2, |p|), 0, 1) with
 *            p = (dx, dy) + 0.5; stores Pixel.floats(b: d, g: 0, r: 0, a: 1) through the raw pointer; `color` unused.
 *   LINE     p = {x0, y0, x1, y1}: float DDA, max(|dx|,|dy|) steps, x and y accumulated in float, `.rounded()`
 *            (to nearest, ties away from zero), bounds-checked store of `color`.
 *   TRIANGLE p = {ax, ay, bx, by, cx, cy}: integer corners, stable sort on int y, the span rule of R:467-494,
 *            colour = the barycentric weights (r = ws.x, g = ws.y, b = ws.z), bounds-checked stores.
 * Apple `simd` pieces defined here (not in the reference tree): length(p) = sqrtf(x*x + y*y) with separately rounded
 * products and sum; smoothstep(e0, e1, x) = t*t*(3 - 2*t), t = clamp((x - e0) / (e1 - e0), 0, 1).
 * Valid domain (the reference corrupts memory or traps outside it): RECT and CIRCLE footprints inside the image,
 * w, h, r != 0 (a zero stride traps); otherwise SR_ORACLE_ERR_DOMAIN and NOTHING is drawn.
 */
enum { SR_ORACLE_2D_RECT = 0, SR_ORACLE_2D_CIRCLE = 1, SR_ORACLE_2D_LINE = 2, SR_ORACLE_2D_TRIANGLE = 3 };
typedef struct sr_oracle_cmd2d {
    int32_t kind;
    int32_t p[6];
    uint8_t color[4]; /* b, g, r, a */
} sr_oracle_cmd2d;
int sr_oracle_draw2d(const sr_oracle_cmd2d* cmds, size_t n, int width, int height, uint8_t* color_bgra);

#ifdef __cplusplus
}
#endif
#endif
