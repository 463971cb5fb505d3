// sr_kernels.cuh — the sm_100a kernels of the rasterization path.  See DESIGN.md for the pipeline.
//
//   k_convert_mesh   (set_mesh)  Swift AoS-32 vertices + int64 indices -> packed xyz / float4 rgb / int32 indices
//   k_transform      K1  Vertex.apply + convertedToScreen per unique vertex            (R:159-172)
//   k_setup          K2  per triangle: sort, truncate, domain rule, screen cull, classify;
//                        small triangles rasterize here with global atomics; large ones are appended
//                        (warp-aggregated) to the bin list and counted per (tile, index-chunk)  (R:238-283)
//   k_scan           K3  single-pass exclusive scan (decoupled look-back) of the (tile, chunk) counters
//   k_bin_fill       K4  writes (triangle -> tile) pairs into per-tile lists, chunks in descending index order
//   k_tile           K5  one CTA per 16x16 tile: row-mask coverage from the integer span rule, visibility
//                        resolve with exact early-out, barycentric shading, fused clear + write-back (R:245-269)
//   k_points_*           PrimitiveType.vertices (R:295-302); also the clear-only path of .line (R:289-293)
#pragma once
#include "sr_math.cuh"

namespace sr {

constexpr int TILE = 16;             // tile edge in pixels
constexpr int TILE_SHIFT = 4;
constexpr int NCHUNK = 64;           // index chunks per tile list (early-out granularity)
constexpr int SMALL_MAX = 8;         // direct path: clipped bbox at most SMALL_MAX x SMALL_MAX pixels
constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 8;        // per thread -> 2048 counters per CTA
constexpr int TILE_BATCH = 16;       // triangles per coverage batch in k_tile

// Per-view counters, read back by the host after the frame.
struct ViewStatus {
    unsigned n_large;        // triangles appended to the bin list
    unsigned n_small;        // triangles rasterized directly
    unsigned domain_skipped; // primitives (or .vertices corners) outside the valid domain
    unsigned total_pairs;    // (triangle, tile) pairs requested (may exceed capacity -> frame is re-run)
};

// Bin-list record of a large triangle: truncated corners in ORIGINAL order (R:251), sort permutation (R:271).
struct __align__(16) TriRec {
    int ax, ay, bx, by, cx, cy;
    unsigned prim;
    unsigned perm;
};

__device__ __forceinline__ SortedTri sorted_of(const TriRec& t) {
    const int xs[3] = {t.ax, t.bx, t.cx};
    const int ys[3] = {t.ay, t.by, t.cy};
    SortedTri s;
    const unsigned i0 = t.perm & 3, i1 = (t.perm >> 2) & 3, i2 = (t.perm >> 4) & 3;
    s.x0 = i0 == 0 ? xs[0] : (i0 == 1 ? xs[1] : xs[2]);  s.y0 = i0 == 0 ? ys[0] : (i0 == 1 ? ys[1] : ys[2]);
    s.x1 = i1 == 0 ? xs[0] : (i1 == 1 ? xs[1] : xs[2]);  s.y1 = i1 == 0 ? ys[0] : (i1 == 1 ? ys[1] : ys[2]);
    s.x2 = i2 == 0 ? xs[0] : (i2 == 1 ? xs[1] : xs[2]);  s.y2 = i2 == 0 ? ys[0] : (i2 == 1 ? ys[1] : ys[2]);
    return s;
}

// ---------------------------------------------------------------------------------------------------------------
// set_mesh: layout conversion + index validation.  first_bad = smallest position of an out-of-range index.
// ---------------------------------------------------------------------------------------------------------------
__global__ void k_convert_vertices(const float4* __restrict__ aos, size_t nverts, float* __restrict__ pos,
                                   float4* __restrict__ col) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nverts) return;
    float4 p = aos[2 * i], c = aos[2 * i + 1];
    pos[3 * i + 0] = p.x; pos[3 * i + 1] = p.y; pos[3 * i + 2] = p.z;
    col[i] = make_float4(c.x, c.y, c.z, 0.0f);
}

__global__ void k_convert_indices(const long long* __restrict__ in, size_t nidx, size_t nverts, int* __restrict__ out,
                                  unsigned long long* __restrict__ first_bad) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nidx) return;
    long long v = in[i];
    bool bad = v < 0 || (unsigned long long)v >= (unsigned long long)nverts;
    out[i] = bad ? 0 : (int)v;
    if (bad) atomicMin(first_bad, (unsigned long long)i);
}

// ---------------------------------------------------------------------------------------------------------------
// K1: vertex transform.  One thread per unique vertex; packed xyz in (12 B), float4 (sx, sy, z, 0) out (16 B).
// The reference transforms per corner (R:224-227); the function is deterministic so per-vertex is identical.
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_transform(const float* __restrict__ pos, int nverts, Mat4 M, float fw, float fh,
                                                   float4* __restrict__ sv) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nverts) return;
    float x = pos[3 * (size_t)i], y = pos[3 * (size_t)i + 1], z = pos[3 * (size_t)i + 2];
    float3 s = transform_to_screen(M, x, y, z, fw, fh);
    sv[i] = make_float4(s.x, s.y, s.z, 0.0f);
}

// ---------------------------------------------------------------------------------------------------------------
// K2: triangle setup, direct raster of small triangles, bin-list append + (tile, chunk) counting of large ones.
// ---------------------------------------------------------------------------------------------------------------
struct FrameParams {
    int width, height;
    int tiles_x, tiles_y;
    int ntris;
    int chunk_shift;          // chunk = prim >> chunk_shift  (< NCHUNK)
    unsigned pair_capacity;
};

template <bool DEPTH>
__device__ __forceinline__ void raster_direct(const SortedTri& s, int ax, int ay, int bx, int by, int cx, int cy,
                                              float za, float zb, float zc, unsigned prim, int y0, int y1, int W,
                                              unsigned* __restrict__ vis32, unsigned long long* __restrict__ vis64) {
    BaryT bt;
    if (DEPTH) bt = bary_setup(ax, ay, bx, by, cx, cy);
    for (int y = y0; y <= y1; ++y) {
        int l, r;
        span_row(s, y, l, r);
        l = max(l, 0);
        r = min(r, W - 1);
        for (int x = l; x <= r; ++x) {
            size_t o = (size_t)y * W + x;
            if (DEPTH) {
                float3 w = bary_weights(bt, x, y);
                float d = blend3(za, zb, zc, w);
                if (depth_can_pass(d)) atomicMin(vis64 + o, depth_key(d, prim));
            } else {
                atomicMax(vis32 + o, prim + 1u);
            }
        }
    }
}

template <bool DEPTH>
__global__ void __launch_bounds__(256) k_setup(const int* __restrict__ idx, const float4* __restrict__ sv, FrameParams fp,
                                               unsigned* __restrict__ vis32, unsigned long long* __restrict__ vis64,
                                               TriRec* __restrict__ recs, unsigned* __restrict__ counts,
                                               ViewStatus* __restrict__ status) {
    const int prim = blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned lane = threadIdx.x & 31;
    bool valid = prim < fp.ntris;
    bool skipped = false, small = false, large = false;
    SortedTri s = {};
    int ax = 0, ay = 0, bx = 0, by = 0, cx = 0, cy = 0, y0 = 0, y1 = -1, xmin = 0, xmax = -1;
    float za = 0, zb = 0, zc = 0;
    unsigned perm = 0;
    if (valid) {
        const int i0 = idx[3 * (size_t)prim], i1 = idx[3 * (size_t)prim + 1], i2 = idx[3 * (size_t)prim + 2];
        const float4 a = sv[i0], b = sv[i1], c = sv[i2];
        if (!(in_domain(a.x) && in_domain(a.y) && in_domain(b.x) && in_domain(b.y) && in_domain(c.x) && in_domain(c.y))) {
            skipped = true;
        } else {
            perm = sort3_perm(a.y, b.y, c.y);                       // R:271 float keys, strict <, stable
            ax = __float2int_rz(a.x); ay = __float2int_rz(a.y);     // simd_long2(...) truncates toward zero
            bx = __float2int_rz(b.x); by = __float2int_rz(b.y);
            cx = __float2int_rz(c.x); cy = __float2int_rz(c.y);
            za = a.z; zb = b.z; zc = c.z;
            TriRec t = {ax, ay, bx, by, cx, cy, (unsigned)prim, perm};
            s = sorted_of(t);
            // result-preserving clip/cull: off-screen pixels are rejected one by one by R:246-250
            y0 = max(s.y0, 0);
            y1 = min(s.y2, fp.height - 1);
            xmin = max(min(ax, min(bx, cx)), 0);
            xmax = min(max(ax, max(bx, cx)), fp.width - 1);
            if (y0 <= y1 && xmin <= xmax) {
                small = (y1 - y0 < SMALL_MAX) && (xmax - xmin < SMALL_MAX);
                large = !small;
            }
        }
    }
    // warp-aggregated counters and bin-list append
    const unsigned m_skip = __ballot_sync(0xffffffffu, skipped);
    const unsigned m_small = __ballot_sync(0xffffffffu, small);
    const unsigned m_large = __ballot_sync(0xffffffffu, large);
    unsigned base = 0;
    if (lane == 0) {
        if (m_skip) atomicAdd(&status->domain_skipped, __popc(m_skip));
        if (m_small) atomicAdd(&status->n_small, __popc(m_small));
        if (m_large) base = atomicAdd(&status->n_large, __popc(m_large));
    }
    base = __shfl_sync(0xffffffffu, base, 0);
    if (large) {
        const unsigned slot = base + __popc(m_large & ((1u << lane) - 1u));
        recs[slot] = TriRec{ax, ay, bx, by, cx, cy, (unsigned)prim, perm};
        // count per (tile, chunk); chunks are stored in DESCENDING index order inside each tile list
        const int cslot = NCHUNK - 1 - (prim >> fp.chunk_shift);
        const int tx0 = xmin >> TILE_SHIFT, tx1 = xmax >> TILE_SHIFT, ty0 = y0 >> TILE_SHIFT, ty1 = y1 >> TILE_SHIFT;
        for (int ty = ty0; ty <= ty1; ++ty)
            for (int tx = tx0; tx <= tx1; ++tx)
                atomicAdd(&counts[(size_t)(ty * fp.tiles_x + tx) * NCHUNK + cslot], 1u);
    }
    if (small) raster_direct<DEPTH>(s, ax, ay, bx, by, cx, cy, za, zb, zc, (unsigned)prim, y0, y1, fp.width, vis32, vis64);
}

// ---------------------------------------------------------------------------------------------------------------
// K3: exclusive scan of the (tile, chunk) counters, single pass with decoupled look-back.
// Writes cursor[i] = exclusive prefix, zeroes counts[i] for the next frame, total -> status->total_pairs.
// Flags carry the launch epoch so the status array never needs clearing.
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(SCAN_THREADS) k_scan(unsigned* __restrict__ counts, unsigned* __restrict__ cursor, int n,
                                                       volatile unsigned long long* __restrict__ look, unsigned* __restrict__ ticket,
                                                       unsigned ticket_base, unsigned epoch, ViewStatus* __restrict__ status) {
    __shared__ unsigned s_bid, s_prefix, s_warp[SCAN_THREADS / 32];
    if (threadIdx.x == 0) s_bid = atomicAdd(ticket, 1u) - ticket_base;
    __syncthreads();
    const unsigned bid = s_bid;
    const int base = (int)bid * SCAN_THREADS * SCAN_ITEMS + threadIdx.x * SCAN_ITEMS;
    unsigned v[SCAN_ITEMS], sum = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) {
        v[k] = (base + k < n) ? counts[base + k] : 0u;
        sum += v[k];
    }
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) if (base + k < n) counts[base + k] = 0u;
    // block exclusive scan of `sum`
    const unsigned lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    unsigned inc = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { unsigned t = __shfl_up_sync(0xffffffffu, inc, d); if (lane >= d) inc += t; }
    if (lane == 31) s_warp[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        unsigned w = lane < SCAN_THREADS / 32 ? s_warp[lane] : 0u, winc = w;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { unsigned t = __shfl_up_sync(0xffffffffu, winc, d); if (lane >= d) winc += t; }
        if (lane < SCAN_THREADS / 32) s_warp[lane] = winc - w;   // exclusive warp offsets
        if (lane == SCAN_THREADS / 32 - 1) {
            const unsigned total = winc;
            // publish aggregate, then look back
            const unsigned long long F_AGG = (unsigned long long)(2u * epoch + 1u) << 32, F_INC = (unsigned long long)(2u * epoch + 2u) << 32;
            unsigned prefix = 0;
            if (bid == 0) {
                look[0] = F_INC | total;
            } else {
                look[bid] = F_AGG | total;
                __threadfence();
                for (int j = (int)bid - 1; j >= 0; --j) {
                    unsigned long long f;
                    do { f = look[j]; } while ((unsigned)(f >> 32) < 2u * epoch + 1u);
                    prefix += (unsigned)f;
                    if ((unsigned)(f >> 32) == 2u * epoch + 2u) break;
                }
                look[bid] = F_INC | (unsigned long long)(prefix + total);
            }
            s_prefix = prefix;
            if ((int)((bid + 1) * SCAN_THREADS * SCAN_ITEMS) >= n) status->total_pairs = prefix + total;
        }
    }
    __syncthreads();
    unsigned run = s_prefix + s_warp[wid] + (inc - sum);
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) {
        if (base + k < n) cursor[base + k] = run;
        run += v[k];
    }
}

// ---------------------------------------------------------------------------------------------------------------
// K4: fill the per-tile lists.  pairs[pos] = bin-list slot.  After this kernel cursor[i] = END of segment i.
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_bin_fill(const TriRec* __restrict__ recs, const ViewStatus* __restrict__ status,
                                                  FrameParams fp, unsigned* __restrict__ cursor, unsigned* __restrict__ pairs) {
    const unsigned n = status->n_large;
    for (unsigned slot = blockIdx.x * blockDim.x + threadIdx.x; slot < n; slot += gridDim.x * blockDim.x) {
        const TriRec t = recs[slot];
        const SortedTri s = sorted_of(t);
        const int y0 = max(s.y0, 0), y1 = min(s.y2, fp.height - 1);
        const int xmin = max(min(t.ax, min(t.bx, t.cx)), 0), xmax = min(max(t.ax, max(t.bx, t.cx)), fp.width - 1);
        const int cslot = NCHUNK - 1 - (int)(t.prim >> fp.chunk_shift);
        const int tx0 = xmin >> TILE_SHIFT, tx1 = xmax >> TILE_SHIFT, ty0 = y0 >> TILE_SHIFT, ty1 = y1 >> TILE_SHIFT;
        for (int ty = ty0; ty <= ty1; ++ty)
            for (int tx = tx0; tx <= tx1; ++tx) {
                const unsigned pos = atomicAdd(&cursor[(size_t)(ty * fp.tiles_x + tx) * NCHUNK + cslot], 1u);
                if (pos < fp.pair_capacity) pairs[pos] = slot;
            }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// K5: tile kernel.  One CTA per 16x16 tile, one thread per pixel for resolve/shade; for coverage the 256 threads
// are (triangle-in-batch, row) pairs that evaluate the integer span rule once per row and publish a 16-bit column
// mask.  Visibility lives in registers:
//   depth off : v = max(prim + 1)            ("highest primitive index wins" = the sequential loop of R:222-229
//                                             with the unconditional store of R:268)
//   depth less: k = min(depth_key(d, prim))  (R:257-261)
// The tile list is ordered by descending index chunk, so with depth off the loop stops as soon as every pixel of
// the tile holds an index no remaining triangle can beat: the image is identical, overdraw work is not done.
// The kernel also consumes (and re-clears) the global visibility words written by the direct path, shades the
// winner (R:251-268) and writes colour and depth exactly once per pixel — the clears of R:205-206 are fused here.
// ---------------------------------------------------------------------------------------------------------------
struct ShadeIn {
    const int* idx;
    const float4* sv;
    const float4* col;
};

__device__ __forceinline__ void shade_pixel(const ShadeIn& in, unsigned prim, int x, int y, bool depth_mode,
                                            unsigned& bgra, float& depth) {
    const int i0 = in.idx[3 * (size_t)prim], i1 = in.idx[3 * (size_t)prim + 1], i2 = in.idx[3 * (size_t)prim + 2];
    const float4 a = in.sv[i0], b = in.sv[i1], c = in.sv[i2];
    const float4 ca = in.col[i0], cb = in.col[i1], cc = in.col[i2];
    const BaryT bt = bary_setup(__float2int_rz(a.x), __float2int_rz(a.y), __float2int_rz(b.x), __float2int_rz(b.y),
                                __float2int_rz(c.x), __float2int_rz(c.y));                      // R:251
    const float3 w = bary_weights(bt, x, y);                                                    // R:252
    if (depth_mode) depth = blend3(a.z, b.z, c.z, w);                                           // R:257
    bgra = pack_bgra(blend3(ca.x, cb.x, cc.x, w), blend3(ca.y, cb.y, cc.y, w), blend3(ca.z, cb.z, cc.z, w));  // R:266-268
}

template <bool DEPTH>
__global__ void __launch_bounds__(256) k_tile(FrameParams fp, ShadeIn in, const TriRec* __restrict__ recs,
                                              const unsigned* __restrict__ cursor, const unsigned* __restrict__ pairs,
                                              unsigned* __restrict__ vis32, unsigned long long* __restrict__ vis64,
                                              unsigned* __restrict__ color_out, float* __restrict__ depth_out) {
    __shared__ __align__(16) unsigned short s_mask[TILE][TILE_BATCH];
    __shared__ unsigned s_prim[TILE_BATCH];
    __shared__ BaryT s_bt[DEPTH ? TILE_BATCH : 1];
    __shared__ float s_z[DEPTH ? TILE_BATCH : 1][3];

    const int tile = blockIdx.x;
    const int tx = tile % fp.tiles_x, ty = tile / fp.tiles_x;
    const int tid = threadIdx.x;
    const int px = tid & (TILE - 1), py = tid >> TILE_SHIFT;
    const int x = tx * TILE + px, y = ty * TILE + py;
    const bool inside = x < fp.width && y < fp.height;
    const size_t o = (size_t)y * fp.width + x;

    unsigned v = 0;
    unsigned long long k = SR_KEY_EMPTY;
    if (inside) {
        if (DEPTH) { k = vis64[o]; if (k != SR_KEY_EMPTY) vis64[o] = SR_KEY_EMPTY; }
        else       { v = vis32[o]; if (v) vis32[o] = 0u; }
    }

    const size_t seg = (size_t)tile * NCHUNK;
    unsigned p = seg == 0 ? 0u : cursor[seg - 1];
    const unsigned end = min(cursor[seg + NCHUNK - 1], fp.pair_capacity);
    const int x_lo = tx * TILE, x_hi = min(tx * TILE + TILE - 1, fp.width - 1);

    while (p < end) {
        const int n = min((unsigned)TILE_BATCH, end - p);
        {   // phase A: (triangle j, row) -> column mask
            const int j = tid >> TILE_SHIFT, row = tid & (TILE - 1);
            unsigned mask = 0;
            if (j < n) {
                const TriRec t = recs[pairs[p + j]];
                const SortedTri s = sorted_of(t);
                const int yy = ty * TILE + row;
                if (yy >= s.y0 && yy <= s.y2 && yy < fp.height) {
                    int l, r;
                    span_row(s, yy, l, r);
                    l = max(l, x_lo);
                    r = min(r, x_hi);
                    if (l <= r) mask = ((2u << (r - l)) - 1u) << (l - x_lo);
                }
                if (row == 0) {
                    s_prim[j] = t.prim;
                    if (DEPTH) {
                        s_bt[j] = bary_setup(t.ax, t.ay, t.bx, t.by, t.cx, t.cy);
                        const int i0 = in.idx[3 * (size_t)t.prim], i1 = in.idx[3 * (size_t)t.prim + 1], i2 = in.idx[3 * (size_t)t.prim + 2];
                        s_z[j][0] = in.sv[i0].z; s_z[j][1] = in.sv[i1].z; s_z[j][2] = in.sv[i2].z;
                    }
                }
            }
            s_mask[row][j] = (unsigned short)mask;
        }
        __syncthreads();
        {   // phase B: each pixel folds the batch into its visibility register
            const uint4 m0 = *reinterpret_cast<const uint4*>(&s_mask[py][0]);
            const uint4 m1 = *reinterpret_cast<const uint4*>(&s_mask[py][8]);
            const unsigned mw[8] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w};
#pragma unroll
            for (int j = 0; j < TILE_BATCH; ++j) {
                const unsigned m = (mw[j >> 1] >> ((j & 1) * 16)) & 0xffffu;
                if (j < n && ((m >> px) & 1u)) {
                    if (DEPTH) {
                        const float3 w = bary_weights(s_bt[j], x, y);
                        const float d = blend3(s_z[j][0], s_z[j][1], s_z[j][2], w);
                        if (depth_can_pass(d)) k = min(k, depth_key(d, s_prim[j]));
                    } else {
                        v = max(v, s_prim[j] + 1u);
                    }
                }
            }
        }
        p += n;
        if (DEPTH) {
            __syncthreads();
        } else {
            // every remaining pair has prim < thr (chunks are stored in descending order)
            const unsigned thr = ((s_prim[n - 1] >> fp.chunk_shift) + 1u) << fp.chunk_shift;
            const bool final_px = !inside || v > thr;    // v - 1 >= thr
            if (__syncthreads_and(final_px)) break;
        }
    }

    if (!inside) return;
    unsigned bgra = 0u;                                  // clear colour (R:205)
    float depth = __int_as_float(0x7f800000);            // +inf (R:206)
    if (DEPTH) { if (k != SR_KEY_EMPTY) shade_pixel(in, (unsigned)k, x, y, true, bgra, depth); }
    else       { if (v) shade_pixel(in, v - 1u, x, y, false, bgra, depth); }
    color_out[o] = bgra;
    if (depth_out) depth_out[o] = depth;
}

// ---------------------------------------------------------------------------------------------------------------
// PrimitiveType.vertices (R:295-302): every corner plots one pixel, the last corner in submission order wins.
// k_points_resolve also serves PrimitiveType.line, which draws nothing (R:289-293): clear only.
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_points_scatter(const int* __restrict__ idx, const float4* __restrict__ sv, int ncorners,
                                                        int W, int H, unsigned* __restrict__ vis32, ViewStatus* __restrict__ status) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    bool skipped = false;
    if (c < ncorners) {
        const float4 s = sv[idx[c]];
        if (!(in_domain(s.x) && in_domain(s.y))) {
            skipped = true;
        } else {
            const int x = __float2int_rz(s.x), y = __float2int_rz(s.y);     // Int(v.xyz.x) R:298-299
            if (x >= 0 && x < W && y >= 0 && y < H) atomicMax(vis32 + (size_t)y * W + x, (unsigned)c + 1u);  // R:30-36
        }
    }
    const unsigned m = __ballot_sync(0xffffffffu, skipped);
    if ((threadIdx.x & 31) == 0 && m) atomicAdd(&status->domain_skipped, __popc(m));
}

__global__ void __launch_bounds__(256) k_points_resolve(const int* __restrict__ idx, const float4* __restrict__ col, size_t npx,
                                                        unsigned* __restrict__ vis32, unsigned* __restrict__ color_out,
                                                        float* __restrict__ depth_out) {
    const size_t o = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (o >= npx) return;
    const unsigned v = vis32[o];
    unsigned bgra = 0u;
    if (v) {
        vis32[o] = 0u;
        const float4 c = col[idx[v - 1u]];
        bgra = pack_bgra(c.x, c.y, c.z);          // Pixel(float3: vertex.color) R:300
    }
    color_out[o] = bgra;
    if (depth_out) depth_out[o] = __int_as_float(0x7f800000);
}

__global__ void k_fill_u64(unsigned long long* __restrict__ p, size_t n, unsigned long long value) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = value;
}

}  // namespace sr
