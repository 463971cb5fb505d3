// sr_oracle.cpp — CPU restatement of Renderer.render(renderPass:).  TEST INFRASTRUCTURE ONLY; see sr_oracle.h.
// "R:" = /root/reference/renderer/Renderer.swift, "A:" = /root/reference/renderer/App.swift.
// Build: g++ -O2 -ffp-contract=off (no -ffast-math) — every float * and + below must round separately,
// as Swift compiles them.
#include "sr_oracle.h"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <limits>

namespace {

struct F3 { float x, y, z; };
struct L2 { int64_t x, y; };          // simd_long2
struct ScreenVertex { F3 xyz; F3 color; };

// ---- simd shim (the three defined operations; header comment items (1)-(3)) ------------------

// simd_muladd(a, b, c): Apple's simd_mul accumulates with it (r = c0*x; r = muladd(c1, y, r); ...).  On x86 without FMA it
// is a*b + c rounded twice (the default here and in csrc/); on arm64 - the platform the reference ships on - the
// compiler contracts it into one fused multiply-add.  -DSR_SIMD_FMA builds that variant (libsr_oracle_fma.so, mirrored
// by libsr_cuda_fma.so), so the uncertainty of "bit-exact coverage" can be measured (tests/test_simd_fma.py).
inline float muladd(float a, float b, float c) {
#ifdef SR_SIMD_FMA
    return fmaf(a, b, c);
#else
    return a * b + c;
#endif
}

// simd_mul(float4x4, float4) for v = (x,y,z,1)  — R:160.  Column accumulate.
inline void mat4_mul_point(const float* m, float x, float y, float z, float out[4]) {
    for (int r = 0; r < 4; ++r) {
        float acc = m[0 * 4 + r] * x;
        acc = muladd(m[1 * 4 + r], y, acc);
        acc = muladd(m[2 * 4 + r], z, acc);
        acc = muladd(m[3 * 4 + r], 1.0f, acc);
        out[r] = acc;
    }
}

struct M2 { float c0x, c0y, c1x, c1y; };  // columns

// matrix_float2x2.inverse — R:99.
inline M2 inverse2x2(const M2& m) {
    float det = m.c0x * m.c1y - m.c1x * m.c0y;
    M2 inv;
    inv.c0x = m.c1y / det;
    inv.c0y = -m.c0y / det;
    inv.c1x = -m.c1x / det;
    inv.c1y = m.c0x / det;
    return inv;
}

inline float clamp01(float v) { return fminf(fmaxf(v, 0.0f), 1.0f); }  // simd_clamp, R:119-122

// ---- Pixel.floats / Pixel(float3:)  R:117-128 ---------------------------------------------------
inline void pack_pixel(F3 c, uint8_t* out) {
    out[0] = (uint8_t)(clamp01(c.z) * 255.0f);  // b
    out[1] = (uint8_t)(clamp01(c.y) * 255.0f);  // g
    out[2] = (uint8_t)(clamp01(c.x) * 255.0f);  // r
    out[3] = (uint8_t)(clamp01(1.0f) * 255.0f); // a
}

// ---- Vertex.apply  R:159-163 and convertedToScreen  R:165-172 -----------------------------------
inline F3 apply_transform(const float* m, F3 p) {
    float v[4];
    mat4_mul_point(m, p.x, p.y, p.z, v);
    return F3{v[0] / v[3], v[1] / v[3], v[2] / v[3]};
}

inline F3 converted_to_screen(F3 p, int width, int height) {
    float ux = p.x * 0.5f + 0.5f;    // R:166
    float uy = p.y * -0.5f + 0.5f;
    // R:168: `.rounded` binds to (Float(W), Float(H)) — a no-op on integers; the product is NOT rounded.
    return F3{ux * (float)width, uy * (float)height, p.z};
}

inline bool in_domain(float s) { return fabsf(s) < 1073741824.0f; }  // finite and |s| < 2^30 (NaN fails)

// ---- interpolate  R:467-494 ----------------------------------------------------------------------
inline int64_t interpolate(const L2* v, int n, int64_t t) {
    int base = 0;
    if (n == 3) {
        if (t >= v[2].y) base = 2;
        else if (t >= v[1].y) base = 1;
    }
    int next = base + 1;
    L2 s = v[base];
    if (next >= n) return s.x;                 // R:478-480
    L2 e = v[next];
    int64_t diff = e.x - s.x;
    int64_t dy = e.y - s.y;
    if (dy == 0) return s.x;                   // R:486-488
    int64_t nt = t - s.y;
    return s.x + diff * nt / dy;               // R:492: left-to-right, truncating division
}

// ---- Triangle.ws / T  R:88-100 ------------------------------------------------------------------
inline M2 triangle_T(L2 a, L2 b, L2 c) {
    float afx = (float)a.x + 0.5f, afy = (float)a.y + 0.5f;
    float bfx = (float)b.x + 0.5f, bfy = (float)b.y + 0.5f;
    float cfx = (float)c.x + 0.5f, cfy = (float)c.y + 0.5f;
    M2 m{afx - cfx, afy - cfy, bfx - cfx, bfy - cfy};
    return inverse2x2(m);
}

inline F3 weights_from_T(const M2& t, float cfx, float cfy, float px, float py) {
    float dx = px - cfx, dy = py - cfy;
    // simd_mul(float2x2, float2): c0*x, then muladd(c1, y, .)  (R:91)
    float wa = muladd(t.c1x, dy, t.c0x * dx);
    float wb = muladd(t.c1y, dy, t.c0y * dx);
    return F3{wa, wb, 1.0f - wa - wb};        // R:92
}

inline F3 triangle_ws(L2 a, L2 b, L2 c, float px, float py) {
    float cfx = (float)c.x + 0.5f, cfy = (float)c.y + 0.5f;   // R:89
    M2 t = triangle_T(a, b, c);                               // R:90
    return weights_from_T(t, cfx, cfy, px, py);
}

inline L2 to_long2(F3 s) { return L2{(int64_t)s.x, (int64_t)s.y}; }  // simd_long2(float2): toward zero

struct Target {
    int width, height;
    uint8_t* color;
    float* depth;
    bool depth_less;
    sr_oracle_stats* st;
};

inline void blend_and_store(const Target& tg, const ScreenVertex& a, const ScreenVertex& b, const ScreenVertex& c,
                            F3 ws, int64_t x, int64_t y) {
    size_t o = (size_t)y * (size_t)tg.width + (size_t)x;     // A:351-360
    if (tg.depth_less) {                                      // R:257-261 uncommented
        float d = a.xyz.z * ws.x + b.xyz.z * ws.y + c.xyz.z * ws.z;
        if (!(d < tg.depth[o])) return;
        tg.depth[o] = d;
        tg.st->depth_passed++;
    }
    F3 col;                                                   // R:266
    col.x = a.color.x * ws.x + b.color.x * ws.y + c.color.x * ws.z;
    col.y = a.color.y * ws.x + b.color.y * ws.y + c.color.y * ws.z;
    col.z = a.color.z * ws.x + b.color.z * ws.y + c.color.z * ws.z;
    pack_pixel(col, tg.color + 4 * o);                        // R:268
}

// ---- draw(triangle:)  R:238-287, exactly as written (unclipped loops, per-pixel T) -------------
void draw_triangle_literal(const Target& tg, const ScreenVertex& a, const ScreenVertex& b, const ScreenVertex& c) {
    // R:271: stable insertion sort on float y with strict <, then truncate.
    const ScreenVertex* s[3] = {&a, &b, &c};
    for (int i = 1; i < 3; ++i)
        for (int j = i; j > 0 && s[j]->xyz.y < s[j - 1]->xyz.y; --j) { const ScreenVertex* t = s[j]; s[j] = s[j - 1]; s[j - 1] = t; }
    L2 sorted[3] = {to_long2(s[0]->xyz), to_long2(s[1]->xyz), to_long2(s[2]->xyz)};
    L2 left[3] = {sorted[0], sorted[1], sorted[2]};           // R:272
    L2 right[2] = {sorted[0], sorted[2]};                     // R:273
    for (int64_t y = sorted[0].y; y <= sorted[2].y; ++y) {    // R:275
        int64_t lx = interpolate(left, 3, y);
        int64_t rx = interpolate(right, 2, y);
        if (lx > rx) { int64_t t = lx; lx = rx; rx = t; }     // R:278-280
        for (int64_t x = lx; x <= rx; ++x) {                  // R:281
            // setPixel R:245-269
            if (!(x >= 0 && x < tg.width && y >= 0 && y < tg.height)) { tg.st->rejected_pixels++; continue; }
            tg.st->span_pixels++;
            F3 ws = triangle_ws(to_long2(a.xyz), to_long2(b.xyz), to_long2(c.xyz), (float)x + 0.5f, (float)y + 0.5f);
            blend_and_store(tg, a, b, c, ws, x, y);
        }
    }
}

// ---- same function with per-triangle invariants hoisted and loops clipped to the image ----------
// Must be bit-identical to the literal form (tests/test_oracle_kat.py checks); this is the timed CPU baseline.
void draw_triangle_hoisted(const Target& tg, const ScreenVertex& a, const ScreenVertex& b, const ScreenVertex& c) {
    const ScreenVertex* s[3] = {&a, &b, &c};
    for (int i = 1; i < 3; ++i)
        for (int j = i; j > 0 && s[j]->xyz.y < s[j - 1]->xyz.y; --j) { const ScreenVertex* t = s[j]; s[j] = s[j - 1]; s[j - 1] = t; }
    L2 left[3] = {to_long2(s[0]->xyz), to_long2(s[1]->xyz), to_long2(s[2]->xyz)};
    L2 right[2] = {left[0], left[2]};
    L2 ia = to_long2(a.xyz), ib = to_long2(b.xyz), ic = to_long2(c.xyz);
    M2 T = triangle_T(ia, ib, ic);
    float cfx = (float)ic.x + 0.5f, cfy = (float)ic.y + 0.5f;
    int64_t y0 = left[0].y < 0 ? 0 : left[0].y;
    int64_t y1 = left[2].y >= tg.height ? tg.height - 1 : left[2].y;
    for (int64_t y = y0; y <= y1; ++y) {
        int64_t lx = interpolate(left, 3, y);
        int64_t rx = interpolate(right, 2, y);
        if (lx > rx) { int64_t t = lx; lx = rx; rx = t; }
        if (lx < 0) lx = 0;
        if (rx >= tg.width) rx = tg.width - 1;
        float py = (float)y + 0.5f;
        for (int64_t x = lx; x <= rx; ++x) {
            tg.st->span_pixels++;
            F3 ws = weights_from_T(T, cfx, cfy, (float)x + 0.5f, py);
            blend_and_store(tg, a, b, c, ws, x, y);
        }
    }
}

}  // namespace

extern "C" {

int sr_oracle_render(const sr_oracle_vertex* vertices, size_t nverts, const int64_t* indices, size_t nidx,
                     const float* transform, int primitive_type, int width, int height,
                     uint8_t* color_bgra, float* depth, uint32_t flags, sr_oracle_stats* stats) {
    sr_oracle_stats local;
    if (!stats) stats = &local;
    std::memset(stats, 0, sizeof(*stats));
    if (width <= 0 || height <= 0 || !color_bgra || !depth || !transform) return SR_ORACLE_ERR_ARG;
    if (primitive_type < SR_ORACLE_TRIANGLE || primitive_type > SR_ORACLE_VERTICES) return SR_ORACLE_ERR_ARG;
    if (nidx && (!indices || !vertices)) return SR_ORACLE_ERR_ARG;

    const int per = primitive_type == SR_ORACLE_LINE ? 2 : 3;      // R:179-188
    const size_t nprims = nidx / per;                              // R:221
    for (size_t i = 0; i < nprims * per; ++i)                      // R:226 would trap
        if (indices[i] < 0 || (uint64_t)indices[i] >= nverts) return SR_ORACLE_ERR_INDEX;

    // clear  R:205-206, R:232-236
    const size_t npx = (size_t)width * (size_t)height;
    std::memset(color_bgra, 0, npx * 4);
    const float inf = std::numeric_limits<float>::infinity();
    for (size_t i = 0; i < npx; ++i) depth[i] = inf;

    Target tg{width, height, color_bgra, depth, (flags & SR_ORACLE_DEPTH_LESS) != 0, stats};
    const bool literal = (flags & SR_ORACLE_LITERAL) != 0;

    for (size_t p = 0; p < nprims; ++p) {                          // R:222
        stats->primitives++;
        if (primitive_type == SR_ORACLE_LINE) continue;            // draw(line:) R:289-293 does nothing
        ScreenVertex sv[3];
        bool ok[3];
        for (int k = 0; k < 3; ++k) {                              // R:224-227, R:241-243 / R:297
            const sr_oracle_vertex& v = vertices[indices[p * 3 + k]];
            F3 ndc = apply_transform(transform, F3{v.x, v.y, v.z});
            sv[k].xyz = converted_to_screen(ndc, width, height);
            sv[k].color = F3{v.r, v.g, v.b};
            ok[k] = in_domain(sv[k].xyz.x) && in_domain(sv[k].xyz.y);
        }
        if (primitive_type == SR_ORACLE_VERTICES) {                // draw(vertices:) R:295-302
            for (int k = 0; k < 3; ++k) {
                if (!ok[k]) { stats->domain_skipped++; continue; }
                int64_t x = (int64_t)sv[k].xyz.x, y = (int64_t)sv[k].xyz.y;
                if (x >= 0 && x < width && y >= 0 && y < height) {  // R:30-36
                    pack_pixel(sv[k].color, color_bgra + 4 * ((size_t)y * width + x));
                    stats->span_pixels++;
                } else {
                    stats->rejected_pixels++;
                }
            }
            continue;
        }
        if (!(ok[0] && ok[1] && ok[2])) { stats->domain_skipped++; continue; }
        if (literal) draw_triangle_literal(tg, sv[0], sv[1], sv[2]);
        else draw_triangle_hoisted(tg, sv[0], sv[1], sv[2]);
    }
    return stats->domain_skipped ? SR_ORACLE_ERR_DOMAIN : SR_ORACLE_OK;
}

void sr_oracle_transform_vertex(const float* transform, const float* xyz, int width, int height, float* out3) {
    F3 s = converted_to_screen(apply_transform(transform, F3{xyz[0], xyz[1], xyz[2]}), width, height);
    out3[0] = s.x; out3[1] = s.y; out3[2] = s.z;
}

int64_t sr_oracle_interpolate(const int64_t* xy, int n, int64_t t) {
    L2 v[3] = {};
    for (int i = 0; i < n && i < 3; ++i) v[i] = L2{xy[2 * i], xy[2 * i + 1]};
    return interpolate(v, n, t);
}

void sr_oracle_weights(const int64_t* abc, int64_t x, int64_t y, float* out3) {
    F3 ws = triangle_ws(L2{abc[0], abc[1]}, L2{abc[2], abc[3]}, L2{abc[4], abc[5]}, (float)x + 0.5f, (float)y + 0.5f);
    out3[0] = ws.x; out3[1] = ws.y; out3[2] = ws.z;
}

void sr_oracle_pack_pixel(const float* rgb, uint8_t* out) { pack_pixel(F3{rgb[0], rgb[1], rgb[2]}, out); }

// ---- legacy 2-D immediate-mode API (R:376-465) -------------------------------------------------------------------
namespace {
inline int sgn(int v) { return (v > 0) - (v < 0); }
inline bool cmd2d_in_domain(const sr_oracle_cmd2d& c, int W, int H) {
    if (c.kind == SR_ORACLE_2D_RECT) {
        const int x = c.p[0], y = c.p[1], w = c.p[2], h = c.p[3];
        if (w == 0 || h == 0) return false;                               // stride(by: 0) traps
        const int ya = y, yb = y + w - sgn(w), xa = x, xb = x + h - sgn(h);   // first / last row and column touched
        return std::min(ya, yb) >= 0 && std::max(ya, yb) < H && std::min(xa, xb) >= 0 && std::max(xa, xb) < W;
    }
    if (c.kind == SR_ORACLE_2D_CIRCLE) {
        const int x = c.p[0], y = c.p[1], r = c.p[2];
        if (r == 0) return false;
        const int a = -r, b = r - sgn(r);
        return y + std::min(a, b) >= 0 && y + std::max(a, b) < H && x + std::min(a, b) >= 0 && x + std::max(a, b) < W;
    }
    return c.kind == SR_ORACLE_2D_LINE || c.kind == SR_ORACLE_2D_TRIANGLE;
}
inline void store_checked(uint8_t* img, int W, int H, int64_t x, int64_t y, const uint8_t* px) {   // Image subscript set, R:30-36
    if (x >= 0 && x < W && y >= 0 && y < H) std::memcpy(img + 4 * ((size_t)y * W + (size_t)x), px, 4);
}
}  // namespace

int sr_oracle_draw2d(const sr_oracle_cmd2d* cmds, size_t n, int W, int H, uint8_t* img) {
    if (!img || W <= 0 || H <= 0 || (n && !cmds)) return SR_ORACLE_ERR_ARG;
    for (size_t k = 0; k < n; ++k) {
        if (cmds[k].kind < SR_ORACLE_2D_RECT || cmds[k].kind > SR_ORACLE_2D_TRIANGLE) return SR_ORACLE_ERR_ARG;
        if (!cmd2d_in_domain(cmds[k], W, H)) return SR_ORACLE_ERR_DOMAIN;
    }
    for (size_t k = 0; k < n; ++k) {
        const sr_oracle_cmd2d& c = cmds[k];
        if (c.kind == SR_ORACLE_2D_RECT) {                                  // R:376-383 (w drives y, h drives x)
            const int x0 = c.p[0], y0 = c.p[1], w = c.p[2], h = c.p[3];
            for (int y = y0; y != y0 + w; y += sgn(w))
                for (int x = x0; x != x0 + h; x += sgn(h)) std::memcpy(img + 4 * ((size_t)y * W + x), c.color, 4);
        } else if (c.kind == SR_ORACLE_2D_CIRCLE) {                         // R:385-403
            const int cx = c.p[0], cy = c.p[1], ri = c.p[2];
            const float r = (float)ri;
            for (int dy = -ri; dy != ri; dy += sgn(ri))
                for (int dx = -ri; dx != ri; dx += sgn(ri)) {
                    const float px = (float)dx + 0.5f, py = (float)dy + 0.5f;
                    const float l = sqrtf(px * px + py * py);
                    float d;
                    if (l < r - 2.0f) {
                        d = 1.0f;
                    } else {
                        const float e0 = r + 2.0f, e1 = r - 2.0f;
                        const float t = clamp01((l - e0) / (e1 - e0));
                        d = clamp01(t * t * (3.0f - 2.0f * t));
                    }
                    uint8_t px4[4] = {(uint8_t)(clamp01(d) * 255.0f), 0, 0, 255};   // .floats(b: d, g: 0, r: 0, a: 1)
                    std::memcpy(img + 4 * ((size_t)(cy + dy) * W + (cx + dx)), px4, 4);
                }
        } else if (c.kind == SR_ORACLE_2D_LINE) {                           // R:405-419
            const int dx = c.p[2] - c.p[0], dy = c.p[3] - c.p[1];
            const int steps = std::max(std::abs(dx), std::abs(dy));
            const float xs = (float)dx / (float)steps, ys = (float)dy / (float)steps;
            float x = (float)c.p[0], y = (float)c.p[1];
            for (int s = 0; s < steps; ++s) {
                store_checked(img, W, H, (int64_t)roundf(x), (int64_t)roundf(y), c.color);
                x += xs;
                y += ys;
            }
        } else {                                                            // R:421-465
            L2 v[3] = {{c.p[0], c.p[1]}, {c.p[2], c.p[3]}, {c.p[4], c.p[5]}};
            L2 s[3] = {v[0], v[1], v[2]};
            for (int i = 1; i < 3; ++i)                                     // stable insertion sort, strict < on int y
                for (int j = i; j > 0 && s[j].y < s[j - 1].y; --j) std::swap(s[j], s[j - 1]);
            const L2 right[2] = {s[0], s[2]};
            for (int64_t y = s[0].y; y <= s[2].y; ++y) {
                int64_t l = interpolate(s, 3, y), r = interpolate(right, 2, y);
                if (l > r) std::swap(l, r);
                if (y < 0 || y >= H) continue;                              // every store of the row fails the bounds check
                for (int64_t x = std::max<int64_t>(l, 0); x <= std::min<int64_t>(r, W - 1); ++x) {
                    const F3 ws = triangle_ws(v[0], v[1], v[2], (float)x + 0.5f, (float)y + 0.5f);
                    // (ac*ws.x + bc*ws.y + cc*ws.z) * a with unit colours and a = 1: each channel is its weight plus zeros
                    const F3 col = {((ws.x * 1.0f + ws.y * 0.0f) + ws.z * 0.0f) * 1.0f, ((ws.x * 0.0f + ws.y * 1.0f) + ws.z * 0.0f) * 1.0f,
                                    ((ws.x * 0.0f + ws.y * 0.0f) + ws.z * 1.0f) * 1.0f};
                    uint8_t px4[4];
                    pack_pixel(col, px4);
                    store_checked(img, W, H, x, y, px4);
                }
            }
        }
    }
    return SR_ORACLE_OK;
}

}  // extern "C"
