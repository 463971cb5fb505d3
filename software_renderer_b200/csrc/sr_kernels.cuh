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
#include <cuda_pipeline.h>

namespace sr {

// Programmatic dependent launch: wait for the previous kernel of the stream (no-op without the launch attribute),
// then let the next one start scheduling its CTAs behind ours.
__device__ __forceinline__ void pdl_wait() {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;");
}

// The span rule's quotient (R:492) in k_tile: 32-bit float estimate + one remainder step (sr_math.cuh lerp_f32), each
// (triangle, row) lane computing its own reciprocals.  Measured against the per-edge FP64 reciprocals it replaced
// (-DSR_DIV_FP64 brings them back): soup1m 6,170 -> 6,289 frames/s, ico80k 17,694 -> 18,379.  Not because FP64 was slow -
// the FP64 pipe carried 2.0 % of k_tile's instructions (profiles/r02_soup1m.md) - but because the reciprocal section ran
// divergently on three lanes of every warp each round and its results went through shared memory.
#if !defined(SR_DIV_FP64) && !defined(SR_DIV32)
#define SR_DIV32 1
#endif
constexpr int TILE = 16;             // tile edge in pixels
constexpr int TILE_SHIFT = 4;
constexpr int NCHUNK = 16;           // index chunks per tile list and batch (early-out granularity)
constexpr int MAX_BATCHES = 8;       // index-descending batches per frame (depth off, heavy overdraw)
constexpr int SMALL_MAX = 8;         // direct path: clipped bbox at most SMALL_MAX x SMALL_MAX pixels
constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 8;        // per thread -> 2048 counters per CTA
constexpr int TILE_BATCH = 16;
constexpr int ZHIST_BINS = 64;
#ifndef SR_SETUP_THREADS
#define SR_SETUP_THREADS 256
#endif       // k_setup CTA size.  Its warps are independent (all staging is per warp; no block barrier after the first
             // instruction).  Measured: 128 threads +0.6 % on the 1M soup (finer release of registers when one warp
             // walks a large triangle), -10 % on the 10M-triangle mesh of pixel-sized triangles (twice the CTA launches
             // for threads that live a few hundred cycles); 64 threads worse on both.
#ifndef SR_SETUP_MINB
#define SR_SETUP_MINB (5 * 256 / SR_SETUP_THREADS)
#endif       // resident CTAs per SM asked of the binning k_setup variants (1280 threads: 47-48 registers)
#ifndef SR_SETUP_DEPTH_MINB
#define SR_SETUP_DEPTH_MINB SR_SETUP_MINB
#endif       // resident CTAs per SM asked of the depth-mode k_setup variants
#ifndef SR_TILE_DEPTH_MINB
#define SR_TILE_DEPTH_MINB 6
#endif       // resident CTAs per SM asked of the depth-mode k_tile.  Measured (soup1m / ico80k depth-LESS frames/s): 4 CTAs (64
             // registers) 2,168 / 9,550; 5 (48 registers, no spill) 2,262 / 10,139; 6 (40 registers, 40 B spill) 2,321 / 10,490
constexpr int SETUP_THREADS = SR_SETUP_THREADS;
#ifndef SR_LATE_PT
#define SR_LATE_PT 2
#endif       // triangles per thread in phase 1 of the compacting (late-batch) k_setup
constexpr int LATE_PT = SR_LATE_PT;
#ifndef SR_LATE_MINB
#define SR_LATE_MINB SR_SETUP_MINB
#endif
#ifndef SR_TILE_MINB
#define SR_TILE_MINB 6
#endif       // resident CTAs per SM asked of the depth-off k_tile (40 registers, 12 bytes of spill).  Measured with the
             // rank-ordered rounds: 5 CTAs 6070 / 16040 frames/s (soup1m / ico80k), 6 CTAs 6190 / 16320, 8 CTAs 6110 / 15380

// Per-view counters, read back by the host after the frame.
struct ViewStatus {
    unsigned n_large[MAX_BATCHES];      // triangles appended to the bin list, per batch
    unsigned total_pairs[MAX_BATCHES];  // (triangle, tile) pairs requested per batch (> capacity -> frame is re-run)
    unsigned n_small;                   // triangles rasterized directly
    unsigned domain_skipped;            // primitives (or .vertices corners) outside the valid domain
    unsigned tiles_done;                // tiles shaded before the final batch (batched mode)
    unsigned batches_run;               // batches that still had unfinished tiles when they started
    unsigned long long pairs_consumed;  // pairs the tile kernel actually processed (early-out feedback)
    unsigned seg_max;                   // slab lists: largest (tile, chunk) segment requested (> slab_cap -> frame is re-run)
    unsigned open_after_first;          // batched mode: tiles still open after the first batch
    unsigned zlo_inv, zhi;              // depth mode: ~min / max of depth_ord(lower bound) over binned triangles
    unsigned long long pairs_culled;    // depth mode: records rejected by their depth bound or never read
    unsigned zhist[ZHIST_BINS];         // two-pass depth frames: histogram of min corner depth over the on-screen triangles
    unsigned back_survivors;            // two-pass depth frames: triangles of the back pass that were not culled
#ifdef SR_TIMELINE
    // debug build: %globaltimer stamps per launch slot (2 * batch + {0 k_setup, 1 k_tile}): first CTA started,
    // first CTA past the grid dependency, last CTA finished.  Gives the overlapped timeline of a frame.
    unsigned long long tl_launch_inv[2 * MAX_BATCHES], tl_go_inv[2 * MAX_BATCHES], tl_end[2 * MAX_BATCHES];
    // per-warp phase times of k_setup (ns since the warp passed the grid dependency), first and later batches:
    // [batch > 0][phase]: 0 corners loaded + classified, 1 staged, 2 binned (large triangles appended), 3 small triangles rastered
    unsigned long long ph_sum[2][8], ph_max[2][8];   // [..][4..6]: compacting k_setup: phase 1 done, its loads arrived, its tests done
    unsigned ph_cnt[2][8];
#endif
};
#ifdef SR_TIMELINE
__device__ __forceinline__ unsigned long long gtime() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
#define TL_BEGIN(st, slot) do { if (threadIdx.x == 0) atomicMax(&(st)->tl_launch_inv[slot], ~gtime()); } while (0)
#define TL_GO(st, slot)    do { if (threadIdx.x == 0) atomicMax(&(st)->tl_go_inv[slot], ~gtime()); } while (0)
#define TL_END(st, slot)   do { if ((threadIdx.x & 31) == 0) atomicMax(&(st)->tl_end[slot], gtime()); } while (0)   /* every warp: thread 0 may finish long before its CTA */
#if SR_TIMELINE + 0 >= 2      /* per-warp phase times: three same-address atomics per warp and phase - they distort a kernel of 27 k warps */
#define TL_PHASE(st, b, k, t0) do { if ((threadIdx.x & 31) == 0) { const unsigned long long d__ = gtime() - (t0); atomicAdd(&(st)->ph_sum[b][k], d__); atomicMax(&(st)->ph_max[b][k], d__); atomicAdd(&(st)->ph_cnt[b][k], 1u); } } while (0)
#else
#define TL_PHASE(st, b, k, t0) do {} while (0)
#endif
#else
#define TL_BEGIN(st, slot) do {} while (0)
#define TL_GO(st, slot)    do {} while (0)
#define TL_END(st, slot)   do {} while (0)
#define TL_PHASE(st, b, k, t0) do {} while (0)
#endif

// Bin-list record of a large triangle (32 B = one sector): corners SORTED as R:271 leaves them and the sort
// permutation (to get back to the original a,b,c order of R:251).  Slab lists hold the records themselves (the tile
// kernel reads them with no indirection); compact lists hold indices into `recs`.  Depth mode keeps the three
// corner depths (original order) in a side array indexed by primitive, ztri[prim].
struct __align__(32) TriRec {
    int x0, y0, x1, y1, x2, y2;
    unsigned prim;
    unsigned perm;                       // sorted position k holds original corner (perm >> 2k) & 3
};

__device__ __forceinline__ SortedTri sort_corners(int ax, int ay, int bx, int by, int cx, int cy, unsigned perm) {
    const unsigned i0 = perm & 3, i1 = (perm >> 2) & 3, i2 = (perm >> 4) & 3;
    SortedTri s;
    s.x0 = i0 == 0 ? ax : (i0 == 1 ? bx : cx);  s.y0 = i0 == 0 ? ay : (i0 == 1 ? by : cy);
    s.x1 = i1 == 0 ? ax : (i1 == 1 ? bx : cx);  s.y1 = i1 == 0 ? ay : (i1 == 1 ? by : cy);
    s.x2 = i2 == 0 ? ax : (i2 == 1 ? bx : cx);  s.y2 = i2 == 0 ? ay : (i2 == 1 ? by : cy);
    return s;
}
// original-order corners (a, b, c) from a record: out[2*orig + {0,1}]
__device__ __forceinline__ void unsort_corners(const TriRec& t, int out[6]) {
    const int xs[3] = {t.x0, t.x1, t.x2}, ys[3] = {t.y0, t.y1, t.y2};
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const unsigned orig = (t.perm >> (2 * k)) & 3;
#pragma unroll
        for (int q = 0; q < 3; ++q) if (orig == (unsigned)q) { out[2 * q] = xs[k]; out[2 * q + 1] = ys[k]; }
    }
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
    pdl_wait();
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nverts) return;
    float x = pos[3 * (size_t)i], y = pos[3 * (size_t)i + 1], z = pos[3 * (size_t)i + 2];
    float3 s = transform_to_screen(M, x, y, z, fw, fh);
    sv[i] = make_float4(s.x, s.y, s.z, 0.0f);
}


// ---------------------------------------------------------------------------------------------------------------
// Frame / batch parameters.
// With the depth test off, visibility is "highest primitive index wins" (R:222-229 + R:268).  A frame can then be
// processed in index-DESCENDING batches: once every pixel of a tile has been covered by a batch, no triangle of a
// later (lower-index) batch can change it, so the tile is shaded and retired (tile_done[tile] = frame epoch) and
// later batches neither bin nor rasterize into it.  The image is identical to the sequential loop's.
// ---------------------------------------------------------------------------------------------------------------
struct FrameParams {
    int width, height;        // the image the kernels work on: the whole frame, or a band of its rows (sr_set_band)
    int y_org;                // first frame row of that image (0 for a whole frame).  Integer corners are moved to
                              //   band-local rows right after the truncation of R:271: the span rule (R:467-494) only
                              //   uses differences of y, so coverage is unchanged; everything that feeds FLOAT math
                              //   (ws / depth, R:251-257) keeps frame coordinates (local row + y_org)
    int tiles_x, tiles_y, ntiles;
    unsigned pair_capacity;
    unsigned epoch;           // frame stamp for tile_done[]
    int batched;              // 1: several index-descending batches (depth off only)
    int slab_cap;             // > 0: tile lists are fixed-capacity slabs, slab_cap entries per (tile, chunk) segment,
                              //      filled directly by k_setup (no k_scan / k_bin_fill); 0: compact lists
    float zthr[NCHUNK - 1];   // depth mode: ascending thresholds; tile-list chunk c holds triangles whose depth lower
                              //   bound is >= zthr[c-1] (chunk 0: everything else), so lists run front to back
    int nobin;                // 1: speculative frame without binning kernels (the previous frame had no large
                              //    triangle); a large triangle is only counted and the host re-runs the frame
    unsigned* done_bits;      // batched frames: one bit per retired tile, `wpr` words per tile row; all zero between
    int wpr;                  //   frames (set by k_tile when it retires a tile, cleared by the final batch's k_tile)
    unsigned* tile_zmax;      // two-pass depth frames: per tile, depth_ord of the farthest pixel after the front pass
                              //   (0xffffffff while a pixel of the tile is still empty)
    unsigned* lbox;           // batched frames: one word per primitive of the later batches = its clipped bounding box in tile
                              //   coordinates (tx0 | tx1 << 8 | ty0 << 16 | ty1 << 24; LBOX_NONE: nothing on screen), written by
                              //   k_late_boxes; nullptr = not used
};
// Late bounding boxes.  In the later batches of a batched frame most triangles find every tile of their bounding box
// retired, but finding that out costs the whole front of the setup: three index loads, nine position loads, three
// transforms with six IEEE divisions.  None of that depends on which tiles are retired.  So a separate kernel,
// k_late_boxes, computes the tile box of every later-batch triangle into lbox[] at full occupancy (37 registers, no
// shared memory, two triangles per thread), and the late k_setup starts from one coalesced word per triangle: everything
// phase 1 needs arrives in one round trip, the bitmap test is branch-free on a shared copy of the bitmap, and only the
// survivors (~5 % on the 1M soup) run the full setup, dealt round-robin to all warps of the CTA.  16.5 + 17.5 us against
// the 39 us of the late k_setup that transforms every triangle itself.
// Tile coordinates must fit a byte (at most 255 tiles per axis: 4080 pixels); larger frames keep the old path.
constexpr unsigned LBOX_NONE = 0xFFFFFFFEu;   // byte 1 (last tile column) = 0xFF cannot occur in a box
constexpr int LBOX_MAX_TILES = 255;
#ifndef SR_PRE_PT
#define SR_PRE_PT 2
#endif       // triangles per thread in the box pre-pass
#ifndef SR_LBOX_PT
#define SR_LBOX_PT 8
#endif       // box words per thread in phase 1 of the late k_setup
constexpr int PRE_PT = SR_PRE_PT, LBOX_PT = SR_LBOX_PT;

// Retired-tile bitmap (batched frames).  It is constant while k_setup runs, so the loads may take the read-only path.
__device__ __forceinline__ bool tile_retired(const FrameParams& fp, int tx, int ty) {
    return (__ldg(&fp.done_bits[ty * fp.wpr + (tx >> 5)]) >> (tx & 31)) & 1u;
}
// the same test on a copy of the bitmap in shared memory (late k_setup with box words: eight tests per thread)
__device__ __forceinline__ bool row_has_open_in(const unsigned* bits, int wpr, int ty, int txa, int txb) {
    const unsigned* row = bits + ty * wpr;
    const int wa = txa >> 5, wb = txb >> 5;
    const unsigned first = 0xffffffffu << (txa & 31), last = 0xffffffffu >> (31 - (txb & 31));
    if (wa == wb) return (~row[wa] & first & last) != 0u;
    if (~row[wa] & first) return true;
    for (int w = wa + 1; w < wb; ++w) if (~row[w]) return true;
    return (~row[wb] & last) != 0u;
}
// any tile of row `ty`, columns [txa, txb], still open?
__device__ __forceinline__ bool row_has_open(const FrameParams& fp, int ty, int txa, int txb) {
    const unsigned* row = fp.done_bits + ty * fp.wpr;
    const int wa = txa >> 5, wb = txb >> 5;
    const unsigned first = 0xffffffffu << (txa & 31), last = 0xffffffffu >> (31 - (txb & 31));
    if (wa == wb) return (~__ldg(&row[wa]) & first & last) != 0u;      // the usual case: the columns sit in one word
    if (~__ldg(&row[wa]) & first) return true;
    for (int w = wa + 1; w < wb; ++w) if (~__ldg(&row[w])) return true;
    return (~__ldg(&row[wb]) & last) != 0u;
}
struct BatchParams {
    int lo, hi;               // primitive range [lo, hi) of this batch
    int chunk_shift;          // chunk = (prim - lo) >> chunk_shift  (< NCHUNK)
    int index;                // batch ordinal
    int final_batch;          // 1: every tile still open is shaded at the end of this batch
    // Two-pass depth frames (SR_FLAG_DEPTH_LESS, heavy overdraw).  zpass 1 = front pass: only triangles whose nearest
    // corner is in front of `zsplit` are set up, binned and rasterized; every tile is shaded, its final keys stay parked
    // in vis64 and its farthest depth goes to tile_zmax.  zpass 2 = back pass: the other triangles, minus every one
    // whose rigorous depth lower bound is behind all the tiles of its bounding box (it cannot pass the strict < of
    // R:258 anywhere); tiles that receive anything are stamped in tile_done and redone by a persistent k_tile from the
    // parked keys.  zpass 0 = the single pass.  The image is the same in all three.
    // Triangles per warp = 32 >> spread_shift.  A batch whose grid does not fill the GPU (the first batch of a batched
    // frame: 125 k triangles = 26 of 64 warp slots per SM, one wave) lasts as long as ONE warp's chain of dependent work,
    // most of it the binning rounds (32 tile-row items per round).  Spreading the triangles over 2x or 4x the warps
    // (every 2nd / 4th lane holds a triangle) shortens that chain; the host picks the shift so that the grid still fits
    // about one wave.  Nothing else changes: every loop below is driven by ballots and prefix sums over the live lanes.
    int spread_shift;
    int zpass;
    float zsplit;
    float zh_lo, zh_scale;    // histogram bin of a depth z: (z - zh_lo) * zh_scale, clamped to [0, ZHIST_BINS)
};

// Spreads `n` items per lane evenly over the warp: calls f(src_lane, item) once per item.  Source-lane data must
// be read from shared memory (f runs divergently).
template <class F>
__device__ __forceinline__ void warp_balance(int n, F&& f) {
    const unsigned lane = threadIdx.x & 31;
    int incl = n;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { int t = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= (unsigned)d) incl += t; }
    const int excl = incl - n;
    const int total = __shfl_sync(0xffffffffu, incl, 31);
    for (int base = 0; base < total; base += 32) {
        const int k = base + (int)lane;
        int j = 0;                                   // largest j with excl[j] <= k
#pragma unroll
        for (int step = 16; step > 0; step >>= 1) {
            const int e = __shfl_sync(0xffffffffu, excl, j + step);
            if (e <= k) j += step;
        }
        const int ej = __shfl_sync(0xffffffffu, excl, j);
        if (k < total) f(j, k - ej);
    }
}

// item -> (item / w, item % w) for 0 <= item < 2^23, 1 <= w: float estimate (off by at most one), exact correction
__device__ __forceinline__ void divmod_tiles(int item, int w, int& q, int& r) {
    q = __float2int_rz(__fmul_rn(__int2float_rn(item), __frcp_rn(__int2float_rn(w))));
    r = item - q * w;
    if (r < 0) { --q; r += w; } else if (r >= w) { ++q; r -= w; }
}

struct __align__(16) StagedTri { int x0, y0, x1, y1, x2, y2; unsigned prim; int ystart; double r01, r12, r02; unsigned perm; unsigned _pad; };   // sorted corners, edge reciprocals, permutation
struct __align__(16) StagedRect { int tx0, ty0, w, slot; };                                   // tile rectangle; w | cslot<<16

// ---------------------------------------------------------------------------------------------------------------
// K2: triangle setup (R:238-243, R:271-273), direct raster of small triangles, bin-list append and per-(tile, chunk)
// counting of large ones.  Both inner loops (rows of small triangles, tiles of large ones) are re-balanced over
// the warp, so one big triangle does not serialize 31 idle lanes.
// ---------------------------------------------------------------------------------------------------------------
// BIN = false: the speculative no-binning frame (fp.nobin): the large-triangle code is compiled out.
// XFORM = true: batched frames transform the three corners here (Vertex.apply + convertedToScreen, R:159-172, the
// same deterministic function k_transform evaluates) and publish them in sv[] for shading, so vertices of triangles
// that no batch ever reaches are never transformed and k_transform is not launched.
struct XformIn {
    const float* pos;
    Mat4 M;
    float fw, fh;
};
// COMPACT = true (later batches of a batched frame, where ~98 % of the triangles find every tile of their bounding box
// retired): phase 1 runs the cheap part only — transform, domain rule, clipped bounding box, retired-tile bitmap — for
// LATE_PT triangles per thread and pushes the survivors' indices on a CTA-wide list; phase 2 runs the full setup on
// the list with dense warps.  Without it every warp that holds ONE survivor walks the whole binning path with 31 idle
// lanes (48 % of the warps at a 2 % survival rate).  Results are identical: visibility is max(index), list order free.
// Clipped bounding box, in tiles, of a triangle with screen corners a, b, c: exactly the tiles the full setup below derives
// (truncation is monotone, so the sorted corners' rows s0.y / s2.y of R:271 are the min / max of the truncated rows: the box
// needs no sort).  Returns the lbox word; `skip` = a corner outside the valid domain (the reference traps, R:251, R:271).
__device__ __forceinline__ unsigned tile_box_word(const FrameParams& fp, float ax, float ay, float bx, float by, float cx, float cy, bool& skip) {
    skip = !(in_domain(ax) && in_domain(ay) && in_domain(bx) && in_domain(by) && in_domain(cx) && in_domain(cy));
    if (skip) return LBOX_NONE;
    const int px0 = __float2int_rz(ax), px1 = __float2int_rz(bx), px2 = __float2int_rz(cx);
    const int qy0 = __float2int_rz(ay), qy1 = __float2int_rz(by), qy2 = __float2int_rz(cy);       // frame rows
    const int ya = max(min(qy0, min(qy1, qy2)) - fp.y_org, 0), yb = min(max(qy0, max(qy1, qy2)) - fp.y_org, fp.height - 1);
    const int xa = max(min(px0, min(px1, px2)), 0), xb = min(max(px0, max(px1, px2)), fp.width - 1);
    if (ya > yb || xa > xb) return LBOX_NONE;
    return (unsigned)(xa >> TILE_SHIFT) | ((unsigned)(xb >> TILE_SHIFT) << 8) | ((unsigned)(ya >> TILE_SHIFT) << 16) | ((unsigned)(yb >> TILE_SHIFT) << 24);
}

// Box pre-pass of a batched frame: lbox[p] for the primitives [0, n) of the later batches.  PRE_PT triangles per thread,
// loads issued together (indices, then positions), the same transform and the same box arithmetic as the full setup.
// A triangle outside the valid domain is counted HERE, once per frame, and never reaches a later batch.
// Ordering (programmatic dependent launch): the kernel follows the first batch's k_setup in the stream.  It lets its
// own dependents be scheduled at once, works WITHOUT waiting for k_setup (it reads only the mesh and the matrix, and
// lbox[] has no reader left: this frame's k_setup CTAs have all passed their own wait, so the previous frame's kernels
// are complete), and waits at its END, so that "k_late_boxes complete" still implies "k_setup complete" for k_tile.
// Its CTAs therefore move in as k_setup's single wave thins out: 150.0 us per frame against 154.2 with the wait at the
// start (1M soup).  (The same work as extra CTAs inside k_setup's own grid faulted, non-deterministically, in a way that
// followed ptxas' register allocation of the fused kernel and no memory access of the source: DESIGN.md §8.)
__global__ void __launch_bounds__(256) k_late_boxes(const int* __restrict__ idx, XformIn xf, FrameParams fp, int n,
                                                    ViewStatus* __restrict__ status) {
    asm volatile("griddepcontrol.launch_dependents;");
    TL_BEGIN(status, 2 * MAX_BATCHES - 2);               // timeline build: the pre-pass shows as "setup7"
    TL_GO(status, 2 * MAX_BATCHES - 2);
    int pid[PRE_PT], ii[PRE_PT][3];
#pragma unroll
    for (int k = 0; k < PRE_PT; ++k) {
        pid[k] = ((int)blockIdx.x * PRE_PT + k) * 256 + (int)threadIdx.x;
        const bool in = pid[k] < n;
#pragma unroll
        for (int q = 0; q < 3; ++q) ii[k][q] = in ? __ldg(&idx[3 * (size_t)pid[k] + q]) : 0;
    }
    float ps[PRE_PT][3][3];
#pragma unroll
    for (int k = 0; k < PRE_PT; ++k)
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            ps[k][q][0] = __ldg(&xf.pos[3 * (size_t)ii[k][q]]); ps[k][q][1] = __ldg(&xf.pos[3 * (size_t)ii[k][q] + 1]); ps[k][q][2] = __ldg(&xf.pos[3 * (size_t)ii[k][q] + 2]);
        }
    unsigned nskip = 0;
#pragma unroll
    for (int k = 0; k < PRE_PT; ++k) {
        float sx[3], sy[3];
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            const float3 t = transform_to_screen<false>(xf.M, ps[k][q][0], ps[k][q][1], ps[k][q][2], xf.fw, xf.fh);
            sx[q] = t.x; sy[q] = t.y;
        }
        bool skip = false;
        const unsigned word = tile_box_word(fp, sx[0], sy[0], sx[1], sy[1], sx[2], sy[2], skip);
        if (pid[k] < n) fp.lbox[pid[k]] = word;
        nskip += __popc(__ballot_sync(0xffffffffu, skip && pid[k] < n));
    }
    if ((threadIdx.x & 31) == 0 && nskip) atomicAdd(&status->domain_skipped, nskip);
    TL_END(status, 2 * MAX_BATCHES - 2);
    asm volatile("griddepcontrol.wait;" ::: "memory");
}

// PRE = true (later batches of a batched frame with fp.lbox; always with COMPACT): phase 1's answer comes from lbox[]
// (see "Late bounding boxes" above).
template <bool DEPTH, bool BIN, bool XFORM, bool COMPACT = false, bool PRE = false>
__global__ void __launch_bounds__(SETUP_THREADS, COMPACT ? SR_LATE_MINB : (DEPTH ? SR_SETUP_DEPTH_MINB : (BIN ? SR_SETUP_MINB : 2048 / SETUP_THREADS))) k_setup(const int* __restrict__ idx, float4* __restrict__ sv, XformIn xf, FrameParams fp,
                                               BatchParams bp, unsigned* __restrict__ vis32,
                                               unsigned long long* __restrict__ vis64, TriRec* __restrict__ recs,
                                               unsigned* __restrict__ counts, TriRec* __restrict__ slab,
                                               float4* __restrict__ ztri, unsigned* __restrict__ tile_done, ViewStatus* __restrict__ status) {
    TL_BEGIN(status, 2 * bp.index);
    pdl_wait();
    TL_GO(status, 2 * bp.index);
#ifdef SR_TIMELINE
    const unsigned long long tl_t0 = gtime();
    const int tl_b = bp.index > 0 ? 1 : 0;
#endif
    __shared__ StagedTri s_tri[SETUP_THREADS];
    __shared__ StagedRect s_rect[BIN ? SETUP_THREADS : 1];
    __shared__ BaryT s_bt[DEPTH ? SETUP_THREADS : 1];
    __shared__ float s_z[DEPTH ? SETUP_THREADS : 1][3];
    __shared__ unsigned s_tally;                            // (warps finished) << 16 | small triangles of the CTA
    __shared__ int s_list[COMPACT ? SETUP_THREADS * (PRE ? LBOX_PT : LATE_PT) : 1];   // COMPACT: primitives that survived phase 1
    __shared__ int s_nsurv;

    const bool late = fp.batched && bp.index > 0;                   // only later batches can meet retired tiles
    // COMPACT && PRE: everything phase 1 needs is requested in ONE round trip - the retired count, the thread's LBOX_PT
    // box words and the CTA's copy of the retired-tile bitmap (at most 8 x 255 words) - before anything is waited for
    constexpr int LBOX_BITS = (COMPACT && PRE) ? 8 * LBOX_MAX_TILES : 1;
    __shared__ unsigned s_bits[LBOX_BITS];
    unsigned w[(COMPACT && PRE) ? LBOX_PT : 1];
    if (COMPACT && PRE) {
        const unsigned retired = late ? *(volatile const unsigned*)&status->tiles_done : 0u;
#pragma unroll
        for (int k = 0; k < LBOX_PT; ++k) {
            const int p = bp.lo + (blockIdx.x * LBOX_PT + k) * SETUP_THREADS + (int)threadIdx.x;
            w[k] = p < bp.hi ? __ldg(&fp.lbox[p]) : LBOX_NONE;
        }
        for (int q = threadIdx.x; q < fp.wpr * fp.tiles_y; q += SETUP_THREADS) s_bits[q] = __ldg(&fp.done_bits[q]);
        if (retired >= (unsigned)fp.ntiles) return;                  // every tile already retired
    } else if (late && status->tiles_done >= (unsigned)fp.ntiles) return;     // every tile already retired
    if (fp.batched && blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&status->batches_run, 1u);
    if (threadIdx.x == 0) { s_tally = 0; s_nsurv = 0; }
    __syncthreads();                                         // all warps arrive together

    const unsigned lane = threadIdx.x & 31;
    const int wbase = threadIdx.x & ~31;
    int nsurv = 1;                                           // !COMPACT: the loop below runs once, one triangle per thread
    if (COMPACT && PRE) {
        // Phase 1 from the pre-pass: LBOX_PT box words per thread (coalesced, all in flight at once), then the retired-tile
        // bitmap; the survivors' indices go on the CTA-wide list (order is free: visibility is max(index)).
        // (the words and the bitmap copy were requested at the top of the kernel; the barrier there published s_bits)
        TL_PHASE(status, tl_b, 5, tl_t0);
        unsigned survbits = 0u;
#pragma unroll
        for (int k = 0; k < LBOX_PT; ++k) {
            bool surv = false;
            if ((w[k] & 0xff00u) != 0xff00u) {
                const int ta = w[k] & 255, tb = (w[k] >> 8) & 255, ra = (w[k] >> 16) & 255, rb = w[k] >> 24;
                // up to eight tile rows, tested without a branch: eight independent shared loads of the word that holds the
                // box's first column (a divergent row-by-row loop with dependent loads cost 2000 cycles per word here, 9 of
                // the kernel's 23 us); a box that crosses a 32-tile word boundary adds its last word, wider ones the loop
                surv = rb - ra >= 8;                        // tall footprints are tested row by row in phase 2
                if (!surv) {
                    const int wa = ta >> 5, wb = tb >> 5;
                    const unsigned first = 0xffffffffu << (ta & 31), last = 0xffffffffu >> (31 - (tb & 31));
                    const unsigned ma = wa == wb ? (first & last) : first;
                    unsigned open = 0u;
#pragma unroll
                    for (int d = 0; d < 8; ++d) open |= ~s_bits[min(ra + d, rb) * fp.wpr + wa] & ma;      // rows past rb repeat rb
                    if (wb != wa) {
#pragma unroll
                        for (int d = 0; d < 8; ++d) open |= ~s_bits[min(ra + d, rb) * fp.wpr + wb] & last;
                        if (wb - wa > 1)
                            for (int ty = ra; ty <= rb && !open; ++ty) open = row_has_open_in(s_bits, fp.wpr, ty, ta, tb) ? 1u : 0u;
                    }
                    surv = open != 0u;
                }
            }
            survbits |= surv ? (1u << k) : 0u;
        }
        // one list reservation per warp for all LBOX_PT words (order inside the list is free)
        unsigned m_surv[LBOX_PT];
        int mine = 0, total = 0;                             // survivors of lower lanes and lower words; of the warp
#pragma unroll
        for (int k = 0; k < LBOX_PT; ++k) {
            m_surv[k] = __ballot_sync(0xffffffffu, (survbits >> k) & 1u);
            total += __popc(m_surv[k]);
        }
        if (total) {                                         // warp-uniform
            int at = 0;
            if (lane == 0) at = atomicAdd(&s_nsurv, total);
            at = __shfl_sync(0xffffffffu, at, 0);
#pragma unroll
            for (int k = 0; k < LBOX_PT; ++k) {
                if ((survbits >> k) & 1u)
                    s_list[at + mine + __popc(m_surv[k] & ((1u << lane) - 1u))] = bp.lo + (blockIdx.x * LBOX_PT + k) * SETUP_THREADS + (int)threadIdx.x;
                mine += __popc(m_surv[k]);
            }
        }
        TL_PHASE(status, tl_b, 6, tl_t0);
        __syncthreads();
        nsurv = s_nsurv;
        TL_PHASE(status, tl_b, 4, tl_t0);
    } else if (COMPACT) {
        // Phase 1.  The loads of the thread's LATE_PT triangles are issued together (indices, then corners): the phase is
        // two dependent memory round trips, so its throughput is the number of loads in flight.
        __shared__ unsigned s_hist[DEPTH ? ZHIST_BINS : 1];
        if (DEPTH && bp.zpass == 1) {
            if (threadIdx.x < ZHIST_BINS) s_hist[threadIdx.x] = 0u;
            __syncthreads();
        }
        unsigned nskip = 0;
        int pid[LATE_PT], ii[LATE_PT][3];
#pragma unroll
        for (int k = 0; k < LATE_PT; ++k) {
            pid[k] = bp.lo + (blockIdx.x * LATE_PT + k) * SETUP_THREADS + (int)threadIdx.x;
            const bool in = pid[k] < bp.hi;
#pragma unroll
            for (int q = 0; q < 3; ++q) ii[k][q] = in ? __ldg(&idx[3 * (size_t)pid[k] + q]) : 0;
        }
        float4 cr[LATE_PT][3];
#pragma unroll
        for (int k = 0; k < LATE_PT; ++k) {
            if (XFORM) {
                float ps[3][3];
#pragma unroll
                for (int q = 0; q < 3; ++q) {
                    ps[q][0] = __ldg(&xf.pos[3 * (size_t)ii[k][q]]); ps[q][1] = __ldg(&xf.pos[3 * (size_t)ii[k][q] + 1]); ps[q][2] = __ldg(&xf.pos[3 * (size_t)ii[k][q] + 2]);
                }
#pragma unroll
                for (int q = 0; q < 3; ++q) {
                    const float3 t = transform_to_screen<false>(xf.M, ps[q][0], ps[q][1], ps[q][2], xf.fw, xf.fh);
                    cr[k][q] = make_float4(t.x, t.y, 0.0f, 0.0f);
                }
            } else {
#pragma unroll
                for (int q = 0; q < 3; ++q) cr[k][q] = __ldg(&sv[ii[k][q]]);
            }
        }
#pragma unroll
        for (int k = 0; k < LATE_PT; ++k) {
            const int p = pid[k];
            const float4 a = cr[k][0], b = cr[k][1], c = cr[k][2];
            bool surv = false, skip = false;
            if (p < bp.hi) {
                if (!(in_domain(a.x) && in_domain(a.y) && in_domain(b.x) && in_domain(b.y) && in_domain(c.x) && in_domain(c.y))) {
                    skip = !(DEPTH && bp.zpass == 2);               // the reference traps here (R:251, R:271); counted once per frame
                } else {
                    // truncation is monotone, so the sorted corners' rows s0.y / s2.y (R:271) are the min / max of the
                    // truncated rows: the clipped bounding box needs no sort
                    const int px0 = __float2int_rz(a.x), px1 = __float2int_rz(b.x), px2 = __float2int_rz(c.x);
                    const int qy0 = __float2int_rz(a.y), qy1 = __float2int_rz(b.y), qy2 = __float2int_rz(c.y);       // frame rows
                    const int ya = max(min(qy0, min(qy1, qy2)) - fp.y_org, 0), yb = min(max(qy0, max(qy1, qy2)) - fp.y_org, fp.height - 1);
                    const int xa = max(min(px0, min(px1, px2)), 0), xb = min(max(px0, max(px1, px2)), fp.width - 1);
                    if (ya <= yb && xa <= xb) {
                        const int ta = xa >> TILE_SHIFT, tb = xb >> TILE_SHIFT, ra = ya >> TILE_SHIFT, rb = yb >> TILE_SHIFT;
                        if (!DEPTH) {
                            surv = !late || rb - ra >= 8;           // tall footprints are tested row by row in phase 2
                            for (int ty = ra; ty <= rb && !surv; ++ty) surv = row_has_open(fp, ty, ta, tb);
                        } else {
                            // front / back by the nearest corner (a NaN or +inf corner depth: fminf skips a NaN; all-NaN compares false: back)
                            const float zmin = fminf(a.z, fminf(b.z, c.z));
                            const bool front = zmin < bp.zsplit;
                            if (bp.zpass == 1) {
                                surv = front;
                                int bin = __float2int_rz(fminf(fmaxf((zmin - bp.zh_lo) * bp.zh_scale, 0.0f), (float)(ZHIST_BINS - 1)));   // NaN -> 0
                                atomicAdd(&s_hist[bin], 1u);
                            } else if (!front) {
                                // behind every tile it can touch?  tile_zmax is what the front pass left (it only gets nearer)
                                unsigned tmax = 0u;
                                if ((tb - ta + 1) * (rb - ra + 1) > 48) tmax = 0xffffffffu;
                                else
                                    for (int ty = ra; ty <= rb; ++ty)
                                        for (int tx = ta; tx <= tb; ++tx) tmax = max(tmax, __ldg(&fp.tile_zmax[ty * fp.tiles_x + tx]));
                                surv = true;
                                // the bound is never above the nearest corner: only then is it worth its four divisions
                                if (tmax != 0xffffffffu && zmin == zmin && depth_ord(zmin) > tmax) {
                                    const float dlo = triangle_depth_floor(px0, qy0, px1, qy1, px2, qy2, a.z, b.z, c.z, xa, ya + fp.y_org, xb, yb + fp.y_org);
                                    surv = !(fabsf(dlo) < __int_as_float(0x7f800000) && depth_ord(dlo) > tmax);
                                }
                            }
                        }
                    }
                }
            }
            const unsigned m_surv = __ballot_sync(0xffffffffu, surv);
            nskip += __popc(__ballot_sync(0xffffffffu, skip));
            if (m_surv) {
                int at = 0;
                if (lane == 0) at = atomicAdd(&s_nsurv, __popc(m_surv));
                at = __shfl_sync(0xffffffffu, at, 0);
                if (surv) s_list[at + __popc(m_surv & ((1u << lane) - 1u))] = p;
            }
        }
        if (lane == 0 && nskip) atomicAdd(&status->domain_skipped, nskip);
        __syncthreads();
        nsurv = s_nsurv;
        if (DEPTH && bp.zpass == 1 && threadIdx.x < ZHIST_BINS && s_hist[threadIdx.x]) atomicAdd(&status->zhist[threadIdx.x], s_hist[threadIdx.x]);
        if (DEPTH && bp.zpass == 2 && threadIdx.x == 0 && nsurv) atomicAdd(&status->back_survivors, (unsigned)nsurv);
    }
    unsigned small_tally = 0;
    for (int lbase = 0; lbase < nsurv; lbase += SETUP_THREADS) {
    // COMPACT: the survivors are dealt round-robin to the CTA's warps (spot = lane * warps + warp), so a short list is
    // shared by all warps instead of filling the first ones: every warp's chain of dependent work (loads, binning
    // rounds, appends) is 1/8 as long
    const int spot = lbase + (int)lane * (SETUP_THREADS / 32) + (int)(threadIdx.x >> 5);
    if (COMPACT && lbase + (int)(threadIdx.x >> 5) >= nsurv) break;            // warp-uniform: this warp has no survivor left
    const int gtid = (int)(blockIdx.x * blockDim.x + threadIdx.x);
    const bool holds = (threadIdx.x & ((1u << bp.spread_shift) - 1u)) == 0u;          // spread: every 2^shift-th lane holds a triangle
    const bool valid = COMPACT ? (spot < nsurv) : (holds && bp.lo + (gtid >> bp.spread_shift) < bp.hi);
    const int prim = COMPACT ? (valid ? s_list[spot] : 0) : bp.lo + (gtid >> bp.spread_shift);
    bool skipped = false, tiny = false, small = false, large = false;
    SortedTri s = {};
    int ax = 0, ay = 0, bx = 0, by = 0, cx = 0, cy = 0, y0 = 0, y1 = -1, xmin = 0, xmax = -1;
    unsigned perm = 0;
    float za = 0.0f, zb = 0.0f, zc = 0.0f;
    float fx[3] = {0.0f, 0.0f, 0.0f}, fy[3] = {0.0f, 0.0f, 0.0f};      // XFORM: screen corners, stored once the triangle survives
    if (valid) {
        const int i0 = idx[3 * (size_t)prim], i1 = idx[3 * (size_t)prim + 1], i2 = idx[3 * (size_t)prim + 2];
        float4 a, b, c;
        if (XFORM) {
            // depth is off in every frame that transforms here, so z_ndc is never read; the corners are stored
            // further down, and only for triangles that can still win a pixel
            auto corner = [&](int i) {
                const float3 t = transform_to_screen<false>(xf.M, xf.pos[3 * (size_t)i], xf.pos[3 * (size_t)i + 1], xf.pos[3 * (size_t)i + 2], xf.fw, xf.fh);
                return make_float4(t.x, t.y, 0.0f, 0.0f);
            };
            a = corner(i0); b = corner(i1); c = corner(i2);
            fx[0] = a.x; fy[0] = a.y; fx[1] = b.x; fy[1] = b.y; fx[2] = c.x; fy[2] = c.y;
        } else {
            a = __ldg(&sv[i0]); b = __ldg(&sv[i1]); c = __ldg(&sv[i2]);     // read-only path (sv is written only under XFORM)
        }
        if (!(in_domain(a.x) && in_domain(a.y) && in_domain(b.x) && in_domain(b.y) && in_domain(c.x) && in_domain(c.y))) {
            skipped = true;                                         // the reference traps here (R:251, R:271)
        } else {
            perm = sort3_perm(a.y, b.y, c.y);                       // R:271 float keys, strict <, stable
            za = a.z; zb = b.z; zc = c.z;
            ax = __float2int_rz(a.x); ay = __float2int_rz(a.y);     // simd_long2(...) truncates toward zero
            bx = __float2int_rz(b.x); by = __float2int_rz(b.y);
            cx = __float2int_rz(c.x); cy = __float2int_rz(c.y);
            s = sort_corners(ax, ay - fp.y_org, bx, by - fp.y_org, cx, cy - fp.y_org, perm);   // band-local rows (|y| < 2^30, no overflow)
            // result-preserving clip/cull: off-screen pixels are rejected one by one by R:246-250
            y0 = max(s.y0, 0);
            y1 = min(s.y2, fp.height - 1);
            xmin = max(min(ax, min(bx, cx)), 0);
            xmax = min(max(ax, max(bx, cx)), fp.width - 1);
            if (y0 <= y1 && xmin <= xmax) {
                // tiny: UNCLIPPED bbox at most 2x2 pixels -> at most three pixels, none of which needs a division
                tiny = (s.y2 - s.y0 <= 1) && (max(ax, max(bx, cx)) - min(ax, min(bx, cx)) <= 1);
                small = !tiny && (y1 - y0 < SMALL_MAX) && (xmax - xmin < SMALL_MAX);
                large = !tiny && !small;
            }
            if (DEPTH && small) {
                s_bt[threadIdx.x] = bary_setup(ax, ay, bx, by, cx, cy);
                s_z[threadIdx.x][0] = a.z; s_z[threadIdx.x][1] = b.z; s_z[threadIdx.x][2] = c.z;
            }
            if (tiny) {
                // rows of R:275 are y0 (and y2 = y0 + 1).  Row y0: left = x2 if y2 == y0 (R:470-471, R:478-480), else
                // x1 if y1 == y0 (nt == 0 on edge s1->s2), else x0; right = x0 (R:486-488 or nt == 0).
                // Row y2 != y0: left = s2.x (R:478-480), right = x0 + diff*dy/dy = s2.x.
                BaryT bt;
                if (DEPTH) bt = bary_setup(ax, ay, bx, by, cx, cy);
                auto plot = [&](int x, int y) {
                    if ((unsigned)x >= (unsigned)fp.width || (unsigned)y >= (unsigned)fp.height) return;   // R:246-250
                    if (late && tile_retired(fp, x >> TILE_SHIFT, y >> TILE_SHIFT)) return;
                    const size_t o = (size_t)y * fp.width + x;
                    if (DEPTH) {
                        if (bp.zpass == 2) tile_done[(y >> TILE_SHIFT) * fp.tiles_x + (x >> TILE_SHIFT)] = fp.epoch;   // back pass: the tile is redone
                        const float d = blend3(a.z, b.z, c.z, bary_weights(bt, x, y + fp.y_org));
                        if (depth_can_pass(d)) atomicMin(vis64 + o, depth_key(d, (unsigned)prim));
                    } else {
                        atomicMax(vis32 + o, (unsigned)prim + 1u);
                    }
                };
                const int l0 = s.y2 == s.y0 ? s.x2 : (s.y1 == s.y0 ? s.x1 : s.x0);
                plot(l0, s.y0);
                if (l0 != s.x0) plot(s.x0, s.y0);
                if (s.y2 != s.y0) plot(s.x2, s.y2);
            }
        }
    }
    int tx0 = xmin >> TILE_SHIFT, tx1 = xmax >> TILE_SHIFT, ty0 = y0 >> TILE_SHIFT, ty1 = y1 >> TILE_SHIFT;
    if (late && (small || large) && !(COMPACT && PRE)) {      // (COMPACT && PRE: phase 1 made this very test, on the same box)
        // retire triangles whose bbox tiles are all done: a few words of the retired-tile bitmap per tile row (taller
        // footprints are tested row by row where their tile-row items are processed)
        if (ty1 - ty0 < 8) {
            bool open = false;
            for (int ty = ty0; ty <= ty1 && !open; ++ty) open = row_has_open(fp, ty, tx0, tx1);
            if (!open) small = large = false;
        }
    }
    if (XFORM && (tiny || small || large)) {     // shading reads the corners of the winner (shared vertices: every
        const int i0 = idx[3 * (size_t)prim], i1 = idx[3 * (size_t)prim + 1], i2 = idx[3 * (size_t)prim + 2];   // writer stores the same bits)
        sv[i0] = make_float4(fx[0], fy[0], 0.0f, 0.0f);
        sv[i1] = make_float4(fx[1], fy[1], 0.0f, 0.0f);
        sv[i2] = make_float4(fx[2], fy[2], 0.0f, 0.0f);
    }
    TL_PHASE(status, tl_b, 0, tl_t0);
    // warp-aggregated counters and bin-list append
    const unsigned m_skip = __ballot_sync(0xffffffffu, skipped);
    const unsigned m_small = __ballot_sync(0xffffffffu, small);
    const unsigned m_large = __ballot_sync(0xffffffffu, large);
    const unsigned m_tiny = __ballot_sync(0xffffffffu, tiny);
    unsigned base = 0;
    if (lane == 0) {
        if (m_skip && !COMPACT) atomicAdd(&status->domain_skipped, __popc(m_skip));     // COMPACT: counted in phase 1
        if (m_large) base = atomicAdd(&status->n_large[bp.index], __popc(m_large));
    }
    const bool binning = BIN && m_large && !fp.nobin;        // warp-uniform
    if (m_small || binning) {
        // reciprocals only for binned triangles (used many times); the few rows of a small one mostly hit the
        // division-free shortcuts of span_row
        if (large && binning)
#ifdef SR_DIV32_SETUP
            s_tri[threadIdx.x] = StagedTri{s.x0, s.y0, s.x1, s.y1, s.x2, s.y2, (unsigned)prim, y0, 0.0, 0.0, 0.0, perm, 0u};
#else
            s_tri[threadIdx.x] = StagedTri{s.x0, s.y0, s.x1, s.y1, s.x2, s.y2, (unsigned)prim, y0,
                                           edge_rcp(s.y1 - s.y0), edge_rcp(s.y2 - s.y1), edge_rcp(s.y2 - s.y0), perm, 0u};
#endif
        else if (small)
            s_tri[threadIdx.x] = StagedTri{s.x0, s.y0, s.x1, s.y1, s.x2, s.y2, (unsigned)prim, y0, 0.0, 0.0, 0.0, perm, 0u};
        __syncwarp();
    }
    // tile row `ty` of staged triangle `src`: the tile columns [ta, tb] that can hold a covered pixel (prunes the
    // corners of the bbox; one evaluation serves the whole tile row)
    auto tile_row_range = [&](int src, const StagedRect& r, int ty, int& ta, int& tb) {
        ta = r.tx0;
        tb = r.tx0 + (r.w & 0x7fff) - 1;
        if (!(r.w & 0x8000)) return;
        const StagedTri t = s_tri[wbase + src];
        const SortedTri st = {t.x0, t.y0, t.x1, t.y1, t.x2, t.y2};
        const int ya = max(max(ty * TILE, t.y0), 0), yb = min(min(ty * TILE + TILE - 1, t.y2), fp.height - 1);
        int mn, mx;
        row_band_extent(st, EdgeRcp{t.r01, t.r12, t.r02}, ya, yb, mn, mx);
        ta = max(ta, mn >> TILE_SHIFT);
        tb = min(tb, mx >> TILE_SHIFT);
    };
    TL_PHASE(status, tl_b, 1, tl_t0);
    if (binning) {
        // compact lists: records are compacted into `recs`, the lists hold their slots; slab lists hold records
        unsigned slot = 0;
        if (DEPTH && large) ztri[prim] = make_float4(za, zb, zc, 0.0f);
        if (!fp.slab_cap) {
            slot = __shfl_sync(0xffffffffu, base, 0) + __popc(m_large & ((1u << lane) - 1u));
            if (large) recs[slot] = TriRec{s.x0, s.y0, s.x1, s.y1, s.x2, s.y2, (unsigned)prim, perm};
        }
        // depth off: chunks are stored in DESCENDING index order inside each tile list
        // depth on : chunks run front to back by a rigorous lower bound of the triangle's depth over its clipped bbox
        int cslot = NCHUNK - 1 - ((prim - bp.lo) >> bp.chunk_shift);
        if (DEPTH && fp.slab_cap) {
            cslot = 0;
            float dlo = __int_as_float(0x7fc00000), zmin = 0.0f, zmax = 0.0f;
            if (large) {
                // the better of two rigorous bounds: over the clipped bbox rectangle, over the covered pixels
                dlo = triangle_depth_floor(ax, ay, bx, by, cx, cy, za, zb, zc, xmin, y0 + fp.y_org, xmax, y1 + fp.y_org);
#pragma unroll
                for (int q = 0; q < NCHUNK - 1; ++q) cslot += (dlo >= fp.zthr[q]) ? 1 : 0;    // NaN: chunk 0
                zmin = fminf(za, fminf(zb, zc)); zmax = fmaxf(za, fmaxf(zb, zc));
            }
            // range of the binned triangles' corner depths: next frame's thresholds
            const bool fin = large && fabsf(zmin) < __int_as_float(0x7f800000) && fabsf(zmax) < __int_as_float(0x7f800000);
            const unsigned hi = __reduce_max_sync(0xffffffffu, fin ? depth_ord(zmax) : 0u);
            const unsigned lo_inv = __reduce_max_sync(0xffffffffu, fin ? ~depth_ord(zmin) : 0u);
            if (lane == 0 && hi) { atomicMax(&status->zhi, hi); atomicMax(&status->zlo_inv, lo_inv); }
        }
        const int items = large ? (tx1 - tx0 + 1) * (ty1 - ty0 + 1) : 0;
        s_rect[threadIdx.x] = StagedRect{tx0, ty0, (tx1 - tx0 + 1) | (items > 2 ? 0x8000 : 0) | (cslot << 16), (int)slot};
        __syncwarp();
        // (a) large triangles: one item per bbox tile ROW, re-balanced over the warp; the item walks the touched tiles.
        //   compact lists: counts[tile][chunk]++ here, k_scan + k_bin_fill build the lists afterwards
        //   slab lists   : the returned count IS the position inside the fixed-capacity segment of (tile, chunk);
        //                  up to four atomics per lane are in flight before the first dependent store
        unsigned seg_max = 0;
        warp_balance(large ? (ty1 - ty0 + 1) : 0, [&](int src, int item) {
            const StagedRect r = s_rect[wbase + src];
            const int cs = r.w >> 16, ty = r.ty0 + item;
            if (late && !((COMPACT && PRE) ? row_has_open_in(s_bits, fp.wpr, ty, r.tx0, r.tx0 + (r.w & 0x7fff) - 1)
                                           : row_has_open(fp, ty, r.tx0, r.tx0 + (r.w & 0x7fff) - 1))) return;
            int ta, tb;
            tile_row_range(src, r, ty, ta, tb);
            const unsigned seg0 = (unsigned)(ty * fp.tiles_x) * NCHUNK + cs;
            if (!fp.slab_cap) {
                for (int tx = ta; tx <= tb; ++tx)
                    if (!(late && tile_retired(fp, tx, ty))) atomicAdd(&counts[seg0 + tx * NCHUNK], 1u);
                return;
            }
            for (int tx = ta; tx <= tb; tx += 4) {
                unsigned pos[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    pos[u] = 0xffffffffu;
                    if (tx + u <= tb && !(late && ((COMPACT && PRE) ? ((s_bits[ty * fp.wpr + ((tx + u) >> 5)] >> ((tx + u) & 31)) & 1u) != 0u : tile_retired(fp, tx + u, ty)))) {
                        pos[u] = atomicAdd(&counts[seg0 + (tx + u) * NCHUNK], 1u);
                        if (DEPTH && bp.zpass == 2) tile_done[ty * fp.tiles_x + tx + u] = fp.epoch;
                    }
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    if (pos[u] == 0xffffffffu) continue;
                    if (pos[u] < (unsigned)fp.slab_cap) {
                        const StagedTri& t = s_tri[wbase + src];
                        slab[(size_t)(seg0 + (tx + u) * NCHUNK) * fp.slab_cap + pos[u]] =
                            TriRec{t.x0, t.y0, t.x1, t.y1, t.x2, t.y2, t.prim, t.perm};
                    } else {
                        seg_max = max(seg_max, pos[u] + 1u);
                    }
                }
            }
        });
        if (seg_max) atomicMax(&status->seg_max, seg_max);          // a segment overflowed: the host grows the slabs and re-runs

    }

    TL_PHASE(status, tl_b, 2, tl_t0);
    // (b) small triangles: one item per clipped row -> global atomics on the visibility words
    if (m_small) {                                           // warp-uniform
        warp_balance(small ? (y1 - y0 + 1) : 0, [&](int src, int item) {
            const StagedTri t = s_tri[wbase + src];
            const SortedTri st = {t.x0, t.y0, t.x1, t.y1, t.x2, t.y2};
            const int y = t.ystart + item;
            int l, r;
            span_row(st, y, l, r);                               // R:276-280
            l = max(l, 0);
            r = min(r, fp.width - 1);
            for (int x = l; x <= r; ++x) {                       // R:281
                if (late && tile_retired(fp, x >> TILE_SHIFT, y >> TILE_SHIFT)) continue;
                const size_t o = (size_t)y * fp.width + x;
                if (DEPTH) {
                    if (bp.zpass == 2 && (x == l || (x & (TILE - 1)) == 0)) tile_done[(y >> TILE_SHIFT) * fp.tiles_x + (x >> TILE_SHIFT)] = fp.epoch;
                    const float3 w = bary_weights(s_bt[wbase + src], x, y + fp.y_org);
                    const float d = blend3(s_z[wbase + src][0], s_z[wbase + src][1], s_z[wbase + src][2], w);
                    if (depth_can_pass(d)) atomicMin(vis64 + o, depth_key(d, t.prim));
                } else {
                    atomicMax(vis32 + o, t.prim + 1u);
                }
            }
        });
    }
    TL_PHASE(status, tl_b, 3, tl_t0);
    small_tally += __popc(m_small) + __popc(m_tiny);
    __syncwarp();                                            // COMPACT: the next pass reuses this warp's staging slots
    }   // survivor passes
    // no barrier at the end (fast warps must not wait for one that bins a screen-sized triangle): the last warp to
    // get here publishes the CTA's small-triangle count
    // (one shared atomic carries both the warp's count and its "done" tick, so no ordering between two words is needed)
    if (lane == 0) {
        const unsigned mine = small_tally;
        const unsigned before = atomicAdd(&s_tally, (1u << 16) | mine);
        if ((before >> 16) == (unsigned)(SETUP_THREADS / 32 - 1)) {
            const unsigned ns = (before & 0xffffu) + mine;
            if (ns) atomicAdd(&status->n_small, ns);
        }
    }
    TL_END(status, 2 * bp.index);
}

// ---------------------------------------------------------------------------------------------------------------
// K3: exclusive scan of the (tile, chunk) counters, single pass with decoupled look-back.
// Writes cursor[i] = exclusive prefix, zeroes counts[i] for the next batch/frame, total -> status->total_pairs[b].
// Flags carry the launch epoch so the look-back array never needs clearing.
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(SCAN_THREADS) k_scan(unsigned* __restrict__ counts, unsigned* __restrict__ cursor, int n,
                                                       volatile unsigned long long* __restrict__ look, unsigned* __restrict__ ticket,
                                                       unsigned ticket_base, unsigned epoch, FrameParams fp, int batch,
                                                       ViewStatus* __restrict__ status) {
    pdl_wait();
    __shared__ unsigned s_bid, s_prefix, s_warp[SCAN_THREADS / 32];
    if (threadIdx.x == 0) s_bid = atomicAdd(ticket, 1u) - ticket_base;
    if (fp.batched && status->tiles_done >= (unsigned)fp.ntiles) return;     // nothing was counted
    __syncthreads();
    const unsigned bid = s_bid;
    const int base = (int)bid * SCAN_THREADS * SCAN_ITEMS + threadIdx.x * SCAN_ITEMS;
    unsigned v[SCAN_ITEMS], sum = 0;
    if (base + SCAN_ITEMS <= n) {
        const uint4 a = *reinterpret_cast<const uint4*>(counts + base), b = *reinterpret_cast<const uint4*>(counts + base + 4);
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
        *reinterpret_cast<uint4*>(counts + base) = make_uint4(0, 0, 0, 0);
        *reinterpret_cast<uint4*>(counts + base + 4) = make_uint4(0, 0, 0, 0);
    } else {
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS; ++k) {
            v[k] = (base + k < n) ? counts[base + k] : 0u;
            if (base + k < n) counts[base + k] = 0u;
        }
    }
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) sum += v[k];
    // block exclusive scan of `sum`
    const unsigned lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    unsigned inc = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { unsigned t = __shfl_up_sync(0xffffffffu, inc, d); if (lane >= (unsigned)d) inc += t; }
    if (lane == 31) s_warp[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        unsigned w = lane < SCAN_THREADS / 32 ? s_warp[lane] : 0u, winc = w;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { unsigned t = __shfl_up_sync(0xffffffffu, winc, d); if (lane >= (unsigned)d) winc += t; }
        if (lane < SCAN_THREADS / 32) s_warp[lane] = winc - w;   // exclusive warp offsets
        const unsigned total = __shfl_sync(0xffffffffu, winc, SCAN_THREADS / 32 - 1);
        const unsigned f_agg = 2u * epoch + 1u, f_inc = 2u * epoch + 2u;
        unsigned prefix = 0;
        if (bid == 0) {
            if (lane == 0) look[0] = ((unsigned long long)f_inc << 32) | total;
        } else {
            if (lane == 0) look[bid] = ((unsigned long long)f_agg << 32) | total;
            // warp-wide look-back: lane l inspects predecessor bid-1-l (32 predecessors per step)
            for (int j = (int)bid - 1 - (int)lane;; j -= 32) {
                unsigned flag = f_inc, val = 0;
                if (j >= 0) {
                    unsigned long long f;
                    do { f = look[j]; } while ((unsigned)(f >> 32) < f_agg);
                    flag = (unsigned)(f >> 32);
                    val = (unsigned)f;
                }
                const unsigned m_inc = __ballot_sync(0xffffffffu, flag == f_inc);
                const int first = m_inc ? __ffs(m_inc) - 1 : 32;         // nearest predecessor with an inclusive prefix
                unsigned add = (int)lane <= first ? val : 0u;
#pragma unroll
                for (int d = 16; d > 0; d >>= 1) add += __shfl_xor_sync(0xffffffffu, add, d);
                prefix += add;
                if (m_inc) break;
            }
            if (lane == 0) look[bid] = ((unsigned long long)f_inc << 32) | (unsigned long long)(prefix + total);
        }
        if (lane == 0) {
            s_prefix = prefix;
            if ((int)((bid + 1) * SCAN_THREADS * SCAN_ITEMS) >= n) status->total_pairs[batch] = prefix + total;
        }
    }
    __syncthreads();
    unsigned run = s_prefix + s_warp[wid] + (inc - sum);
    unsigned o[SCAN_ITEMS];
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) { o[k] = run; run += v[k]; }
    if (base + SCAN_ITEMS <= n) {
        *reinterpret_cast<uint4*>(cursor + base) = make_uint4(o[0], o[1], o[2], o[3]);
        *reinterpret_cast<uint4*>(cursor + base + 4) = make_uint4(o[4], o[5], o[6], o[7]);
    } else {
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS; ++k) if (base + k < n) cursor[base + k] = o[k];
    }
}

// ---------------------------------------------------------------------------------------------------------------
// K4: fill the per-tile lists.  pairs[pos] = bin-list slot.  After this kernel cursor[i] = END of segment i.
// One record per lane, the (record, tile) pairs re-balanced over the warp.
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_bin_fill(const TriRec* __restrict__ recs, const ViewStatus* __restrict__ status,
                                                  FrameParams fp, BatchParams bp, unsigned* __restrict__ cursor,
                                                  unsigned* __restrict__ pairs, const unsigned* __restrict__ tile_done) {
    pdl_wait();
    __shared__ StagedRect s_rect[256];
    if (fp.batched && status->tiles_done >= (unsigned)fp.ntiles) return;
    const unsigned n = status->n_large[bp.index];
    const int wbase = threadIdx.x & ~31;
    const unsigned n_round = (n + 31u) & ~31u;          // whole warps stay in the loop (shuffles inside)
    for (unsigned slot = blockIdx.x * blockDim.x + threadIdx.x; slot < n_round; slot += gridDim.x * blockDim.x) {
        int items = 0;
        if (slot < n) {
            const TriRec t = recs[slot];
            const int y0 = max(t.y0, 0), y1 = min(t.y2, fp.height - 1);
            const int xmin = max(min(t.x0, min(t.x1, t.x2)), 0), xmax = min(max(t.x0, max(t.x1, t.x2)), fp.width - 1);
            const int cslot = NCHUNK - 1 - (((int)t.prim - bp.lo) >> bp.chunk_shift);
            const int tx0 = xmin >> TILE_SHIFT, tx1 = xmax >> TILE_SHIFT, ty0 = y0 >> TILE_SHIFT, ty1 = y1 >> TILE_SHIFT;
            items = (tx1 - tx0 + 1) * (ty1 - ty0 + 1);
            s_rect[threadIdx.x] = StagedRect{tx0, ty0, (tx1 - tx0 + 1) | (items > 2 ? 0x8000 : 0) | (cslot << 16), (int)slot};
        }
        __syncwarp();
        // (record, tile) pairs spread over the warp, four per lane per step so that four list-cursor atomics are in
        // flight before the first dependent store
        const unsigned lane = threadIdx.x & 31;
        int incl = items;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { int t = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= (unsigned)d) incl += t; }
        const int excl = incl - items;
        const int total = __shfl_sync(0xffffffffu, incl, 31);
        for (int base = 0; base < total; base += 128) {
            unsigned pos[4], val[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int k = base + u * 32 + (int)lane;
                int j = 0;
#pragma unroll
                for (int step = 16; step > 0; step >>= 1) {
                    const int e = __shfl_sync(0xffffffffu, excl, j + step);
                    if (e <= k) j += step;
                }
                const int ej = __shfl_sync(0xffffffffu, excl, j);
                pos[u] = 0xffffffffu;
                if (k < total) {
                    const StagedRect r = s_rect[wbase + j];
                    const int item = k - ej, w = r.w & 0x7fff, cs = r.w >> 16;
                    int qy, qx;
                    divmod_tiles(item, w, qy, qx);
                    const int tile = (r.ty0 + qy) * fp.tiles_x + r.tx0 + qx;
                    val[u] = (unsigned)r.slot;
                    bool keep = !(fp.batched && tile_done[tile] == fp.epoch);
                    if (keep && (r.w & 0x8000)) {             // same pruning rule as the counting pass in k_setup
                        const TriRec& t = recs[r.slot];
                        const SortedTri st = {t.x0, t.y0, t.x1, t.y1, t.x2, t.y2};
                        const int tx = r.tx0 + qx, ty = r.ty0 + qy;
                        const int ya = max(max(ty * TILE, t.y0), 0), yb = min(min(ty * TILE + TILE - 1, t.y2), fp.height - 1);
                        // tile COLUMNS of the band's extent, exactly as k_setup counted them (a partial last column
                        // is counted even when the extent starts right of the image): count == fill, no list gaps
                        int mn, mx;
                        row_band_extent(st, EdgeRcp{edge_rcp(t.y1 - t.y0), edge_rcp(t.y2 - t.y1), edge_rcp(t.y2 - t.y0)}, ya, yb, mn, mx);
                        keep = (mn >> TILE_SHIFT) <= tx && tx <= (mx >> TILE_SHIFT);
                    }
                    if (keep)
                        pos[u] = atomicAdd(&cursor[(size_t)tile * NCHUNK + cs], 1u);
                }
            }
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (pos[u] < fp.pair_capacity) pairs[pos[u]] = val[u];
        }
        __syncwarp();
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
// In batched mode a tile that is not yet fully covered parks its visibility words in global memory for the next
// batch; a fully covered one is shaded and retired.
// ---------------------------------------------------------------------------------------------------------------
struct ShadeIn {
    const int* idx;
    const float4* sv;
    const float4* col;
    const float4* ztri;     // depth mode: corner depths of binned triangles, by primitive
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

template <bool DEPTH, bool LATE>          // LATE: a later batch of a batched frame (persistent grid, retired tiles are skipped)
__global__ void __launch_bounds__(256, DEPTH ? SR_TILE_DEPTH_MINB : SR_TILE_MINB) k_tile(FrameParams fp, BatchParams bp, ShadeIn in, const TriRec* __restrict__ recs,
                                              const unsigned* __restrict__ cursor, unsigned* __restrict__ counts,
                                              const unsigned* __restrict__ pairs, unsigned* __restrict__ vis32, unsigned long long* __restrict__ vis64,
                                              unsigned* __restrict__ tile_done, ViewStatus* __restrict__ status,
                                              unsigned* __restrict__ color_out, float* __restrict__ depth_out) {
    TL_BEGIN(status, 2 * bp.index + 1);
    pdl_wait();
    TL_GO(status, 2 * bp.index + 1);
    __shared__ __align__(16) unsigned short s_mask[TILE][TILE_BATCH];
    __shared__ unsigned s_prim[TILE_BATCH];
    __shared__ TriRec s_rec[2][TILE_BATCH];
#ifndef SR_DIV32
    __shared__ double s_rcp[TILE_BATCH][3];                 // per-edge reciprocals of the current round's triangles (-DSR_DIV_FP64)
#endif
    __shared__ BaryT s_bt[DEPTH ? TILE_BATCH : 1];
    __shared__ float s_z[DEPTH ? TILE_BATCH : 1][3];
#ifdef SR_VEC_WB
    __shared__ __align__(16) unsigned s_outc[TILE * TILE];   // experiment: tile staged for 16-byte row-segment stores
    __shared__ __align__(16) float s_outd[TILE * TILE];
#endif
    __shared__ unsigned s_off[NCHUNK + 1];                 // slab lists: prefix of the tile's 16 segment lengths
    __shared__ __align__(16) unsigned s_wmax[8];           // depth mode: per-warp max of the pixels' depth_ord
    __shared__ unsigned s_zord[NCHUNK];                    // depth mode: depth_ord of the chunk thresholds
    __shared__ unsigned s_dlo[DEPTH ? TILE_BATCH : 1];     // depth mode: depth_ord of each round triangle's lower bound over this tile (0 = none)

    const int tid = threadIdx.x;
    const int px = tid & (TILE - 1), py = tid >> TILE_SHIFT;
    const int j = tid >> TILE_SHIFT, row = tid & (TILE - 1);        // coverage role: (triangle j of the batch, row)
    constexpr bool late = LATE;                                     // only later batches can meet retired tiles
    if (DEPTH) {
#pragma unroll
        for (int q = 0; q < NCHUNK - 1; ++q) if (tid == q) s_zord[q] = depth_ord(fp.zthr[q]);
    }
    if (fp.batched && bp.final_batch)       // nothing reads the retired-tile bitmap any more: leave it zero for the next frame
        for (int w = blockIdx.x * blockDim.x + tid; w < fp.wpr * fp.tiles_y; w += gridDim.x * blockDim.x) fp.done_bits[w] = 0u;
    if (late && status->tiles_done >= (unsigned)fp.ntiles) { TL_END(status, 2 * bp.index + 1); return; }

    auto do_tile = [&](const int tile) {
    int tx, ty;
    divmod_tiles(tile, fp.tiles_x, ty, tx);
    const int x = tx * TILE + px, y = ty * TILE + py;
    const bool inside = x < fp.width && y < fp.height;
    const size_t o = (size_t)y * fp.width + x;

    // visibility words left by the direct path (or parked by an earlier batch); re-armed after the loop
    unsigned v = 0;
    unsigned long long k = SR_KEY_EMPTY;
    if (inside) {
        if (DEPTH) k = vis64[o];
        else       v = vis32[o];
    }
    const unsigned v_in = v;
    const unsigned long long k_in = k;
    // depth mode: farthest current depth over the tile's pixels (empty pixel = 0xffffffff = nothing can be rejected)
    auto publish_zmax = [&]() {
        const unsigned wm = __reduce_max_sync(0xffffffffu, inside ? (unsigned)(k >> 32) : 0u);
        if ((tid & 31) == 0) s_wmax[tid >> 5] = wm;
    };
    auto read_zmax = [&]() -> unsigned {
        const uint4 a = *reinterpret_cast<const uint4*>(&s_wmax[0]), b = *reinterpret_cast<const uint4*>(&s_wmax[4]);
        return max(max(max(a.x, a.y), max(a.z, a.w)), max(max(b.x, b.y), max(b.z, b.w)));
    };

    const size_t seg = (size_t)tile * NCHUNK;
    const bool slab = fp.slab_cap > 0 && !fp.nobin;
    unsigned begin = 0u, end = 0u;                         // nobin: no tile lists exist
    if (!slab && !fp.nobin) {
        begin = seg == 0 ? 0u : cursor[seg - 1];
        end = min(cursor[seg + NCHUNK - 1], fp.pair_capacity);
    }
    const int x_lo = tx * TILE, x_hi = min(tx * TILE + TILE - 1, fp.width - 1);
    const int yy = ty * TILE + row;

    __syncthreads();                                       // previous tile of this CTA is done with shared memory
    if (DEPTH) publish_zmax();
    if (slab) {
        // the tile's list = its 16 segments, highest index chunk first; lengths are read and re-armed here
        if (tid < 32) {
            unsigned c = 0;
            if (tid < NCHUNK) {
                c = counts[seg + tid];
                if (c) counts[seg + tid] = 0u;
                c = min(c, (unsigned)fp.slab_cap);         // an overflowed segment: the frame is re-run anyway
            }
            unsigned inc = c;
#pragma unroll
            for (int d = 1; d < NCHUNK; d <<= 1) { unsigned t = __shfl_up_sync(0xffffffffu, inc, d); if (tid >= d) inc += t; }
            if (tid < NCHUNK) s_off[tid + 1] = inc;
            if (tid == 0) s_off[0] = 0u;
        }
        __syncthreads();
        end = s_off[NCHUNK];
        if (tid == 0 && end) atomicAdd(&status->total_pairs[bp.index], end);
    }
    unsigned p = begin;
    unsigned n_culled = 0;

    // software pipeline run by warp 0: lanes 0..15 bring the records of the NEXT round global -> shared with cp.async
    // (no registers held) while the current round is processed; the block barrier that ends the round publishes them.
    const bool producer = tid < 32;
    auto fetch = [&](int buf, unsigned q) {                // lanes 0..15 of warp 0, q = list position of triangle `tid`
        const TriRec* src;
        if (slab) {
            int c = 0;
#pragma unroll
            for (int step = NCHUNK / 2; step > 0; step >>= 1) if (s_off[c + step] <= q) c += step;
            src = recs + (seg + c) * (size_t)fp.slab_cap + (q - s_off[c]);
        } else {
            src = recs + pairs[q];
        }
        __pipeline_memcpy_async(reinterpret_cast<char*>(&s_rec[buf][tid]), reinterpret_cast<const char*>(src), 16);
        __pipeline_memcpy_async(reinterpret_cast<char*>(&s_rec[buf][tid]) + 16, reinterpret_cast<const char*>(src) + 16, 16);
    };
    if (producer) {
        if (tid < TILE_BATCH && p + tid < end) fetch(0, p + tid);
        __pipeline_commit();
        __pipeline_wait_prior(0);
    }
    int cur = 0;
    bool all_final = false;                                // depth off: the list was left early, every pixel holds its final index
    __syncthreads();                                       // s_rec[0] complete

    while (p < end) {
        const int n = min((unsigned)TILE_BATCH, end - p);
        if (producer) {
            if (tid < TILE_BATCH && p + TILE_BATCH + tid < end) fetch(cur ^ 1, p + TILE_BATCH + tid);
            __pipeline_commit();
        }
        unsigned zmax_u = 0xffffffffu;
        if (DEPTH) {
            zmax_u = read_zmax();
            // front-to-back lists: every record from position p on sits in chunk >= c, whose triangles cannot go
            // below zthr[c-1]; once that is beyond the farthest pixel of the tile, nothing left can pass R:258
            if (slab) {
                int c = 0;
#pragma unroll
                for (int step = NCHUNK / 2; step > 0; step >>= 1) if (s_off[c + step] <= p) c += step;
                if (c > 0 && s_zord[c - 1] > zmax_u) break;
            }
        }
        // per-edge reciprocals of this round's triangles: rows 0..2 of a triangle's 16 lanes compute one each
#ifndef SR_DIV32
        if (j < n && row < 3) {
            const TriRec& t = s_rec[cur][j];
            s_rcp[j][row] = edge_rcp(row == 0 ? t.y1 - t.y0 : (row == 1 ? t.y2 - t.y1 : t.y2 - t.y0));
        }
#endif
        if (DEPTH && j < n && row == 0) {
            const TriRec& t = s_rec[cur][j];
            int oc[6];
            unsort_corners(t, oc);
            s_bt[j] = bary_setup(oc[0], oc[1] + fp.y_org, oc[2], oc[3] + fp.y_org, oc[4], oc[5] + fp.y_org);   // frame rows (R:251)
            const float4 z = in.ztri[t.prim];
            s_z[j][0] = z.x; s_z[j][1] = z.y; s_z[j][2] = z.z;
        }
        __syncwarp();
        bool culled = false;
        if (DEPTH) {
            // reject the whole (triangle, tile) record when a rigorous lower bound of its depth over the tile's pixels
            // is beyond the farthest pixel: each lane samples one corner pixel of the tile, xor-shuffles reduce
            const int jq = j < n ? j : 0;
            DepthSample m = depth_sample(s_bt[jq], s_z[jq][0], s_z[jq][1], s_z[jq][2],
                                         (row & 1) ? x_hi : x_lo, ((row & 2) ? min(ty * TILE + TILE - 1, fp.height - 1) : ty * TILE) + fp.y_org);
#pragma unroll
            for (int d = 1; d <= 2; d <<= 1) {
                m.d = fminf(m.d, __shfl_xor_sync(0xffffffffu, m.d, d));
                m.a = fmaxf(m.a, __shfl_xor_sync(0xffffffffu, m.a, d));
                m.b = fmaxf(m.b, __shfl_xor_sync(0xffffffffu, m.b, d));
            }
            // a NaN sample needs an inf/NaN constant of the triangle, which makes the margin non-finite at some corner:
            // then dlo is not finite and nothing is rejected
            const float dlo = depth_lower_bound(m.d, m.a, m.b, s_z[jq][0], s_z[jq][1], s_z[jq][2]);
            const bool bounded = fabsf(dlo) < __int_as_float(0x7f800000);
            culled = bounded && depth_ord(dlo) > zmax_u;
            if (j < n && row == 0) s_dlo[j] = bounded ? depth_ord(dlo) : 0u;      // phase B: a pixel already nearer than this skips the triangle
        }
        {   // phase A: (triangle j, row) -> column mask
            unsigned mask = 0;
            if (j < n && !culled) {
                const TriRec& t = s_rec[cur][j];
                const SortedTri s = {t.x0, t.y0, t.x1, t.y1, t.x2, t.y2};
                if (yy >= s.y0 && yy <= s.y2 && yy < fp.height) {
                    int l, r;
#ifdef SR_DIV32
                    span_row_f32(s, yy, l, r);              // experiment: 32-bit float quotient + one remainder step, no per-edge reciprocals
#else
                    span_row_rcp(s, EdgeRcp{s_rcp[j][0], s_rcp[j][1], s_rcp[j][2]}, yy, l, r);
#endif
                    l = max(l, x_lo);
                    r = min(r, x_hi);
                    if (l <= r) mask = ((2u << (r - l)) - 1u) << (l - x_lo);
                }
            }
            // Depth off: the round's triangles are filed by DESCENDING index (rank = how many of the round have a
            // higher index; a tile list holds a triangle once, so ranks are a permutation), in the column whose bit
            // comes out at position `rank` of the gathered word below: the winner of a pixel is then its lowest set
            // bit, one step instead of a loop over every covering triangle (max is order independent: same image).
            int col = j;
            if (!DEPTH) {
                const unsigned mine = s_rec[cur][j].prim;
                const unsigned higher = __ballot_sync(0xffffffffu, row < n && s_rec[cur][row].prim > mine);
                const int rank = j < n ? __popc((higher >> (tid & 16)) & 0xffffu) : j;      // j >= n: the unused ranks
                col = ((rank & 7) << 1) | (rank >> 3);     // ranks 0..7 -> bits 0..7, ranks 8..15 -> bits 16..23
            }
            if (j < n && row == 0) {
                s_prim[col] = s_rec[cur][j].prim;
                if (DEPTH && culled) ++n_culled;
            }
            s_mask[row][col] = (unsigned short)mask;       // zero for j >= n
        }
        __syncthreads();
        {   // phase B: each pixel gathers "which triangles of the batch cover me" into one word (two triangles per
            // shift/and: bit k = triangle 2k, bit 16+k = triangle 2k+1), then folds only the set bits
            const uint4 m0 = *reinterpret_cast<const uint4*>(&s_mask[py][0]);
            const uint4 m1 = *reinterpret_cast<const uint4*>(&s_mask[py][8]);
            const unsigned mw[8] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w};
            unsigned bits = 0;
#pragma unroll
            for (int q = 0; q < 8; ++q) bits |= ((mw[q] >> px) & 0x00010001u) << q;
            if (!DEPTH) {
                if (bits) {                                // lowest set bit = highest index of the round (see phase A)
                    const int b = __ffs(bits) - 1;
                    v = max(v, s_prim[((b & 15) << 1) | (b >> 4)] + 1u);
                }
            } else {
                while (bits) {
                    const int b = __ffs(bits) - 1;
                    bits &= bits - 1;
                    const int jj = ((b & 15) << 1) | (b >> 4);
                    if (s_dlo[jj] > (unsigned)(k >> 32)) continue;           // every depth of this triangle in the tile is behind the pixel (strict < fails)
                    const float3 w = bary_weights(s_bt[jj], x, y + fp.y_org);
                    const float d = blend3(s_z[jj][0], s_z[jj][1], s_z[jj][2], w);
                    if (depth_can_pass(d)) k = min(k, depth_key(d, s_prim[jj]));
                }
            }
        }
        p += n;
        cur ^= 1;
        if (producer) __pipeline_wait_prior(0);            // next records landed (made visible by the barrier below)
        if (DEPTH) {
            publish_zmax();
            __syncthreads();
        } else {
            // every remaining pair of this batch has prim < thr (chunks are stored in descending order)
            // (the round's lowest index, rank n - 1, sits in the chunk of the round's last record: chunks are monotone in the index)
            const unsigned low = s_prim[(((n - 1) & 7) << 1) | ((n - 1) >> 3)];
            const unsigned thr = (unsigned)bp.lo + ((((low - (unsigned)bp.lo) >> bp.chunk_shift) + 1u) << bp.chunk_shift);
            const bool final_px = !inside || v > thr;    // v - 1 >= thr
            if (__syncthreads_and(final_px)) { all_final = true; break; }
        }
    }
    __pipeline_wait_prior(0);
    if (tid == 0 && p > begin) atomicAdd(&status->pairs_consumed, (unsigned long long)(p - begin));
    if (DEPTH) {
        // records rejected by their bound (counted by the row-0 lanes) + records never read
        unsigned tot = n_culled;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, d);
        if ((tid & 31) == 0 && tot) atomicAdd(&status->pairs_culled, (unsigned long long)tot);
        if (tid == 0 && end > p) atomicAdd(&status->pairs_culled, (unsigned long long)(end - p));
    }

    bool shade_now = true;
    if (fp.batched && !bp.final_batch) {
        // a pixel covered by this or an earlier batch cannot be changed by later (lower-index) batches
        shade_now = all_final || __syncthreads_and(!inside || v != 0u) != 0;    // all_final (block-uniform): v > thr >= 0 everywhere
        if (shade_now && tid == 0) {
            tile_done[tile] = fp.epoch;
            atomicOr(&fp.done_bits[ty * fp.wpr + (tx >> 5)], 1u << (tx & 31));
            atomicAdd(&status->tiles_done, 1u);
        }
    }
    if (DEPTH && bp.zpass == 1) {
        // front pass of a two-pass depth frame: the tile's farthest depth is what the back pass is culled against
        __syncthreads();                                   // every warp is done reading s_wmax
        publish_zmax();
        __syncthreads();
        if (tid == 0) fp.tile_zmax[tile] = read_zmax();
    }
#ifdef SR_VEC_WB
    // experiment: interior tiles are staged in shared memory and leave as 16-byte row segments
    // (64 threads cover a plane) instead of one 4-byte store per thread and plane
    if (!shade_now) { if (inside && v != v_in) vis32[o] = v; return; }      // block-uniform: park for the next batch
    unsigned bgra = 0u;                                  // clear colour (R:205)
    float depth = __int_as_float(0x7f800000);            // +inf (R:206)
    if (inside) {
        if (DEPTH) {
            if (bp.zpass == 1) { if (k != k_in) vis64[o] = k; }
            else if (k_in != SR_KEY_EMPTY) vis64[o] = SR_KEY_EMPTY;
        } else if (v_in) vis32[o] = 0u;
        if (DEPTH) { if (k != SR_KEY_EMPTY) shade_pixel(in, (unsigned)k, x, y + fp.y_org, true, bgra, depth); }
        else       { if (v) shade_pixel(in, v - 1u, x, y + fp.y_org, false, bgra, depth); }
    }
    const bool vec = tx * TILE + TILE <= fp.width && ty * TILE + TILE <= fp.height && (fp.width & 3) == 0 &&
                     (reinterpret_cast<uintptr_t>(color_out) & 15) == 0 && (reinterpret_cast<uintptr_t>(depth_out) & 15) == 0;
    if (vec) {
        s_outc[tid] = bgra;
        s_outd[tid] = depth;
        __syncthreads();
        const int t2 = tid & 63, r = t2 >> 2, sg = t2 & 3;
        const size_t off = (size_t)(ty * TILE + r) * fp.width + tx * TILE + sg * 4;
        if (tid < 64) *reinterpret_cast<uint4*>(color_out + off) = *reinterpret_cast<const uint4*>(&s_outc[r * TILE + sg * 4]);
        else if (tid < 128 && depth_out) *reinterpret_cast<float4*>(depth_out + off) = *reinterpret_cast<const float4*>(&s_outd[r * TILE + sg * 4]);
    } else if (inside) {
        color_out[o] = bgra;
        if (depth_out) depth_out[o] = depth;
    }
#else
    if (!inside) return;
    if (!shade_now) { if (v != v_in) vis32[o] = v; return; }      // park for the next batch
    if (DEPTH) {
        if (bp.zpass == 1) { if (k != k_in) vis64[o] = k; }       // keys stay parked for the back pass (its k_tile re-arms them)
        else if (k_in != SR_KEY_EMPTY) vis64[o] = SR_KEY_EMPTY;   // re-arm the visibility word for the next frame
    }
    else       { if (v_in) vis32[o] = 0u; }
    unsigned bgra = 0u;                                  // clear colour (R:205)
    float depth = __int_as_float(0x7f800000);            // +inf (R:206)
    if (DEPTH) { if (k != SR_KEY_EMPTY) shade_pixel(in, (unsigned)k, x, y + fp.y_org, true, bgra, depth); }
    else       { if (v) shade_pixel(in, v - 1u, x, y + fp.y_org, false, bgra, depth); }
    color_out[o] = bgra;
    if (depth_out) depth_out[o] = depth;
#endif
    };   // do_tile
    if (!LATE) {                                            // first (or only) batch: the grid has one CTA per tile
        do_tile(blockIdx.x);
    } else {
        // persistent CTAs: a late batch that finds most tiles retired costs a few flag reads, not thousands of CTA
        // launches; the flags of a CTA's next 32 tiles are fetched by warp 0 in ONE round trip (one dependent load per
        // tile made the almost empty final batch of the 1M soup a chain of ~10 round trips: 70 % of its stall samples)
        // (depth mode, back pass of a two-pass frame: the tiles to do are those the back pass stamped; every other tile
        // only has its parked keys re-armed)
        __shared__ unsigned s_open, s_clean;
        for (int g0 = blockIdx.x; g0 < fp.ntiles; g0 += 32 * (int)gridDim.x) {
            __syncthreads();                               // s_open of the previous group has been read by everyone
            if (tid < 32) {
                const int t = g0 + tid * (int)gridDim.x;
                const bool stamped = t < fp.ntiles && tile_done[t] == fp.epoch;
                const unsigned m = __ballot_sync(0xffffffffu, t < fp.ntiles && (DEPTH ? stamped : !stamped));   // depth off: retired by an earlier batch?
                const unsigned cl = __ballot_sync(0xffffffffu, DEPTH && t < fp.ntiles && !stamped);
                if (tid == 0) { s_open = m; s_clean = cl; }
            }
            __syncthreads();
            if (DEPTH) {
                unsigned clean = s_clean;
                while (clean) {
                    const int tile = g0 + (__ffs(clean) - 1) * (int)gridDim.x;
                    clean &= clean - 1u;
                    int tx, ty;
                    divmod_tiles(tile, fp.tiles_x, ty, tx);
                    const int x = tx * TILE + px, y = ty * TILE + py;
                    if (x < fp.width && y < fp.height) vis64[(size_t)y * fp.width + x] = SR_KEY_EMPTY;
                }
            }
            unsigned open_mask = s_open;
            while (open_mask) {
                const int tile = g0 + (__ffs(open_mask) - 1) * (int)gridDim.x;
                open_mask &= open_mask - 1u;
                do_tile(tile);
            }
        }
    }
    TL_END(status, 2 * bp.index + 1);
}

// ---------------------------------------------------------------------------------------------------------------
// K5': resolve kernel of the no-binning frames (fp.nobin: every triangle went through the direct path, so the
// visibility words are final).  One thread per four consecutive pixels: 16-B loads of the words, shade the winners
// (R:251-268), 16-B stores of colour and depth (clear fused, R:205-206), words re-armed for the next frame.
// ---------------------------------------------------------------------------------------------------------------
template <bool DEPTH, int PPT>                              // PPT pixels per thread: 4 (large images) or 1 (enough threads)
__global__ void __launch_bounds__(256) k_resolve(ShadeIn in, int W, int y_org, size_t npx, unsigned* __restrict__ vis32,
                                                 unsigned long long* __restrict__ vis64, unsigned* __restrict__ color_out,
                                                 float* __restrict__ depth_out) {
    pdl_wait();
    const size_t o = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * PPT;
    if (o >= npx) return;
    const int cnt = PPT == 1 ? 1 : (int)min((size_t)4, npx - o);
    unsigned win[4] = {0u, 0u, 0u, 0u};                     // prim + 1, 0 = background
    if (cnt == 4) {
        if (DEPTH) {
            const ulonglong2 a = *reinterpret_cast<const ulonglong2*>(vis64 + o), b = *reinterpret_cast<const ulonglong2*>(vis64 + o + 2);
            const unsigned long long k[4] = {a.x, a.y, b.x, b.y};
            bool any = false;
#pragma unroll
            for (int q = 0; q < 4; ++q) if (k[q] != SR_KEY_EMPTY) { win[q] = (unsigned)k[q] + 1u; any = true; }
            if (any) {
                const ulonglong2 e = make_ulonglong2(SR_KEY_EMPTY, SR_KEY_EMPTY);
                *reinterpret_cast<ulonglong2*>(vis64 + o) = e;
                *reinterpret_cast<ulonglong2*>(vis64 + o + 2) = e;
            }
        } else {
            const uint4 v = *reinterpret_cast<const uint4*>(vis32 + o);
            win[0] = v.x; win[1] = v.y; win[2] = v.z; win[3] = v.w;
            if (v.x | v.y | v.z | v.w) *reinterpret_cast<uint4*>(vis32 + o) = make_uint4(0u, 0u, 0u, 0u);
        }
    } else {
        for (int q = 0; q < cnt; ++q) {
            if (DEPTH) { const unsigned long long k = vis64[o + q]; if (k != SR_KEY_EMPTY) { win[q] = (unsigned)k + 1u; vis64[o + q] = SR_KEY_EMPTY; } }
            else       { win[q] = vis32[o + q]; if (win[q]) vis32[o + q] = 0u; }
        }
    }
    int y = (int)(o / (size_t)W), x = (int)(o - (size_t)y * W);
    y += y_org;                                             // shading works in frame rows
    unsigned bgra[4];
    float depth[4];
#pragma unroll
    for (int q = 0; q < PPT; ++q) {
        bgra[q] = 0u;                                       // clear colour (R:205)
        depth[q] = __int_as_float(0x7f800000);              // +inf (R:206)
        if (win[q]) shade_pixel(in, win[q] - 1u, x, y, DEPTH, bgra[q], depth[q]);
        if (++x == W) { x = 0; ++y; }
    }
    if (cnt == 4) {
        *reinterpret_cast<uint4*>(color_out + o) = make_uint4(bgra[0], bgra[1], bgra[2], bgra[3]);
        if (depth_out) *reinterpret_cast<float4*>(depth_out + o) = make_float4(depth[0], depth[1], depth[2], depth[3]);
    } else {
        for (int q = 0; q < cnt; ++q) { color_out[o + q] = bgra[q]; if (depth_out) depth_out[o + q] = depth[q]; }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// PrimitiveType.vertices (R:295-302): every corner plots one pixel, the last corner in submission order wins.
// k_points_resolve also serves PrimitiveType.line, which draws nothing (R:289-293): clear only.
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_points_scatter(const int* __restrict__ idx, const float4* __restrict__ sv, int ncorners,
                                                        int W, int H, int y_org, unsigned* __restrict__ vis32, ViewStatus* __restrict__ status) {
    pdl_wait();
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    bool skipped = false;
    if (c < ncorners) {
        const float4 s = sv[idx[c]];
        if (!(in_domain(s.x) && in_domain(s.y))) {
            skipped = true;
        } else {
            const int x = __float2int_rz(s.x), y = __float2int_rz(s.y) - y_org;     // Int(v.xyz.x) R:298-299; band-local row
            if (x >= 0 && x < W && y >= 0 && y < H) atomicMax(vis32 + (size_t)y * W + x, (unsigned)c + 1u);  // R:30-36
        }
    }
    const unsigned m = __ballot_sync(0xffffffffu, skipped);
    if ((threadIdx.x & 31) == 0 && m) atomicAdd(&status->domain_skipped, __popc(m));
}

__global__ void __launch_bounds__(256) k_points_resolve(const int* __restrict__ idx, const float4* __restrict__ col, size_t npx,
                                                        unsigned* __restrict__ vis32, unsigned* __restrict__ color_out,
                                                        float* __restrict__ depth_out) {
    pdl_wait();
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

// ---------------------------------------------------------------------------------------------------------------
// Legacy 2-D immediate-mode API (R:376-465).  One kernel per command, launched in command order on one stream.
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ int sgn_i(int v) { return (v > 0) - (v < 0); }

// draw(rect:) R:376-383: rect.w drives y, rect.h drives x (as written); unchecked stores.
__global__ void __launch_bounds__(256) k2d_rect(unsigned* __restrict__ img, int W, int x0, int y0, int w, int h, unsigned color) {
    const int nx = abs(h), ny = abs(w);
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nx * ny) return;
    const int iy = t / nx, ix = t - iy * nx;
    img[(size_t)(y0 + iy * sgn_i(w)) * W + (x0 + ix * sgn_i(h))] = color;
}

// draw(circle:) R:385-403: length = sqrt(x*x + y*y) and smoothstep = t*t*(3 - 2t), every operation rounded separately
// (the definition the test oracle shares, see include/sr_cuda.h); only the blue channel carries d.
__global__ void __launch_bounds__(256) k2d_circle(unsigned* __restrict__ img, int W, int cx, int cy, int ri) {
    const int n = 2 * abs(ri);
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n * n) return;
    const int iy = t / n, ix = t - iy * n;
    const int dy = -ri + iy * sgn_i(ri), dx = -ri + ix * sgn_i(ri);
    const float r = __int2float_rn(ri);
    const float px = __fadd_rn(__int2float_rn(dx), 0.5f), py = __fadd_rn(__int2float_rn(dy), 0.5f);
    const float l = __fsqrt_rn(__fadd_rn(__fmul_rn(px, px), __fmul_rn(py, py)));
    float d;
    if (l < __fsub_rn(r, 2.0f)) {
        d = 1.0f;
    } else {
        const float e0 = __fadd_rn(r, 2.0f), e1 = __fsub_rn(r, 2.0f);
        const float u = fminf(fmaxf(__fdiv_rn(__fsub_rn(l, e0), __fsub_rn(e1, e0)), 0.0f), 1.0f);
        d = fminf(fmaxf(__fmul_rn(__fmul_rn(u, u), __fsub_rn(3.0f, __fmul_rn(2.0f, u))), 0.0f), 1.0f);
    }
    const unsigned b = (unsigned)__float2int_rz(__fmul_rn(fminf(fmaxf(d, 0.0f), 1.0f), 255.0f));
    img[(size_t)(cy + dy) * W + (cx + dx)] = b | 0xFF000000u;              // .floats(b: d, g: 0, r: 0, a: 1)
}

// draw(line:) R:405-419: the float accumulation x += xStep is sequential by definition: one thread walks the line.
__global__ void k2d_line(unsigned* __restrict__ img, int W, int H, int x0, int y0, int x1, int y1, unsigned color) {
    if (blockIdx.x | threadIdx.x) return;
    const int dx = x1 - x0, dy = y1 - y0;
    const int steps = max(abs(dx), abs(dy));
    const float xs = __fdiv_rn(__int2float_rn(dx), __int2float_rn(steps)), ys = __fdiv_rn(__int2float_rn(dy), __int2float_rn(steps));
    float x = __int2float_rn(x0), y = __int2float_rn(y0);
    for (int s = 0; s < steps; ++s) {
        const long long px = (long long)roundf(x), py = (long long)roundf(y);   // .rounded(): to nearest, ties away from zero
        if (px >= 0 && px < W && py >= 0 && py < H) img[(size_t)py * W + (size_t)px] = color;   // bounds-checked setter R:30-36
        x = __fadd_rn(x, xs);
        y = __fadd_rn(y, ys);
    }
}

// draw(triangle:with:in:) R:421-465: one thread per row of the span rule; colour = barycentric weights.
__global__ void __launch_bounds__(128) k2d_triangle(unsigned* __restrict__ img, int W, int H, int ax, int ay, int bx, int by, int cx, int cy) {
    // stable sort on INT y (R:422), strict <
    int sx[3] = {ax, bx, cx}, sy[3] = {ay, by, cy};
#pragma unroll
    for (int i = 1; i < 3; ++i)
        for (int j = i; j > 0 && sy[j] < sy[j - 1]; --j) {
            int t = sy[j]; sy[j] = sy[j - 1]; sy[j - 1] = t;
            t = sx[j]; sx[j] = sx[j - 1]; sx[j - 1] = t;
        }
    const SortedTri s = {sx[0], sy[0], sx[1], sy[1], sx[2], sy[2]};
    const long long row = (long long)max(s.y0, 0) + (long long)blockIdx.x * blockDim.x + threadIdx.x;   // rows outside the image store nothing
    if (row > (long long)min(s.y2, H - 1)) return;
    const int y = (int)row;
    int l, r;
    span_row(s, y, l, r);
    l = max(l, 0);
    r = min(r, W - 1);
    const BaryT bt = bary_setup(ax, ay, bx, by, cx, cy);
    for (int x = l; x <= r; ++x) {
        const float3 w = bary_weights(bt, x, y);
        // (ac*ws.x + bc*ws.y + cc*ws.z) * a with unit colours, a = 1 (R:448-452)
        const float cr = __fmul_rn(__fadd_rn(__fadd_rn(__fmul_rn(w.x, 1.0f), __fmul_rn(w.y, 0.0f)), __fmul_rn(w.z, 0.0f)), 1.0f);
        const float cg = __fmul_rn(__fadd_rn(__fadd_rn(__fmul_rn(w.x, 0.0f), __fmul_rn(w.y, 1.0f)), __fmul_rn(w.z, 0.0f)), 1.0f);
        const float cb = __fmul_rn(__fadd_rn(__fadd_rn(__fmul_rn(w.x, 0.0f), __fmul_rn(w.y, 0.0f)), __fmul_rn(w.z, 1.0f)), 1.0f);
        img[(size_t)y * W + x] = pack_bgra(cr, cg, cb);
    }
}

// The depth clear of R:206 as its own kernel (host API, depth test off: the pass never writes depth again, so the
// cleared image can travel to the host while the frame renders).
__global__ void __launch_bounds__(256) k_clear_depth(float* __restrict__ p, size_t n) {
    const float inf = __int_as_float(0x7f800000);
    const size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i + 4 <= n) *reinterpret_cast<float4*>(p + i) = make_float4(inf, inf, inf, inf);
    else for (size_t k = i; k < n; ++k) p[k] = inf;
}

__global__ void k_fill_u64(unsigned long long* __restrict__ p, size_t n, unsigned long long value) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = value;
}

}  // namespace sr
