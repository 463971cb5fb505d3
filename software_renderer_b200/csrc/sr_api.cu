// sr_api.cu — host side of libsr_cuda.so: context, buffers, launch sequence, C ABI (include/sr_cuda.h).
// There is no CPU fallback anywhere in this file: without a CUDA device sr_create fails.
#include "../../include/sr_cuda.h"
#include "sr_kernels.cuh"

#include <algorithm>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include <dlfcn.h>
#if defined(__SSE2__)
#include <emmintrin.h>
#endif

using namespace sr;

namespace {

constexpr int MAX_PROF = 64;
constexpr size_t kSlabLimit = 8ull << 30;

inline bool poison_new_buffers() {
    const char* e = getenv("SR_POISON");     // read per allocation (rare), so a test can switch it on for one context
    return e && atoi(e) != 0;
}

template <class T>
struct DevBuf {
    T* p = nullptr;
    size_t cap = 0;   // elements
    cudaError_t reserve(size_t n) {
        if (n <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        cudaError_t e = cudaMalloc((void**)&p, n * sizeof(T));
        if (e == cudaSuccess) cap = n;
        if (e == cudaSuccess && poison_new_buffers()) {      // SR_POISON=1 (tests): a read of never-written memory faults
            e = cudaMemset(p, 0xFF, n * sizeof(T));
            if (e == cudaSuccess) e = cudaDeviceSynchronize();
        }
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};

// Helper threads of a context for the host-side halves of the host-pointer calls: writing the constant +inf depth
// plane of a pass that never tests depth (R:206; instead of shipping 4 bytes per pixel of a constant over PCIe) and
// moving pageable caller memory through page-locked bounce buffers.  One job at a time, split evenly over the threads.
class HostPool {
public:
    enum Kind { FILL_INF, COPY };
    explicit HostPool(int n) : n_(n) {
        for (int i = 0; i < n_; ++i) th_.emplace_back([this, i] { run(i); });
    }
    ~HostPool() {
        { std::lock_guard<std::mutex> l(m_); stop_ = true; }
        cv_job_.notify_all();
        for (auto& t : th_) t.join();
    }
    int threads() const { return n_; }
    void start(Kind kind, void* dst, const void* src, size_t bytes) {
        if (n_ == 0) { part(Job{kind, dst, src, bytes}, 0, 1); return; }
        {
            std::unique_lock<std::mutex> l(m_);
            cv_done_.wait(l, [&] { return pending_ == 0; });
            job_ = Job{kind, dst, src, bytes};
            pending_ = n_;
            ++gen_;
        }
        cv_job_.notify_all();
    }
    void wait() {
        if (n_ == 0) return;
        std::unique_lock<std::mutex> l(m_);
        cv_done_.wait(l, [&] { return pending_ == 0; });
    }
private:
    struct Job { Kind kind; void* dst; const void* src; size_t bytes; };
    static void part(const Job& j, int id, int n) {
        // 64-byte aligned slices of [0, bytes)
        const size_t per = ((j.bytes + n - 1) / n + 63) & ~(size_t)63;
        const size_t lo = std::min(j.bytes, per * id), hi = std::min(j.bytes, per * (id + 1));
        if (hi <= lo) return;
        char* d = static_cast<char*>(j.dst) + lo;
        if (j.kind == COPY) { std::memcpy(d, static_cast<const char*>(j.src) + lo, hi - lo); return; }
        // +inf floats (R:206).  Streaming stores: the plane is written once and read by somebody else.
        float* f = reinterpret_cast<float*>(d);
        size_t cnt = (hi - lo) / sizeof(float), k = 0;
        const float inf = __builtin_inff();
#if defined(__SSE2__)
        while (k < cnt && (reinterpret_cast<uintptr_t>(f + k) & 15)) f[k++] = inf;
        const __m128 v = _mm_set1_ps(inf);
        for (; k + 16 <= cnt; k += 16) {
            _mm_stream_ps(f + k, v); _mm_stream_ps(f + k + 4, v); _mm_stream_ps(f + k + 8, v); _mm_stream_ps(f + k + 12, v);
        }
        _mm_sfence();
#endif
        for (; k < cnt; ++k) f[k] = inf;
    }
    void run(int id) {
        unsigned long long seen = 0;
        for (;;) {
            Job j;
            {
                std::unique_lock<std::mutex> l(m_);
                cv_job_.wait(l, [&] { return stop_ || gen_ != seen; });
                if (stop_) return;
                seen = gen_;
                j = job_;
            }
            part(j, id, n_);
            {
                std::lock_guard<std::mutex> l(m_);
                if (--pending_ == 0) cv_done_.notify_all();
            }
        }
    }
    int n_;
    std::vector<std::thread> th_;
    std::mutex m_;
    std::condition_variable cv_job_, cv_done_;
    Job job_{FILL_INF, nullptr, nullptr, 0};
    unsigned long long gen_ = 0;
    int pending_ = 0;
    bool stop_ = false;
};

// NCCL, bound at run time (sr_gather): the library the process already holds (PyTorch's bundled copy) or the system's.
struct Nccl {
    typedef struct { char internal[128]; } UniqueId;
    typedef void* Comm;
    int (*GetUniqueId)(UniqueId*) = nullptr;
    int (*CommInitRank)(Comm*, int, UniqueId, int) = nullptr;
    int (*CommDestroy)(Comm) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    int (*Send)(const void*, size_t, int /*ncclDataType_t*/, int, Comm, cudaStream_t) = nullptr;
    int (*Recv)(void*, size_t, int, int, Comm, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    bool ok = false;
    std::string why;
    static Nccl& get() {
        static Nccl n;
        static std::once_flag once;
        std::call_once(once, [] {
            void* h = nullptr;
            const char* names[] = {getenv("SR_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
            for (const char* nm : names) if (nm && !h) h = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
            if (!h) { n.why = "libnccl.so.2 not found (set SR_NCCL_LIB)"; return; }
            auto sym = [&](const char* s) { void* p = dlsym(h, s); if (!p) n.why = std::string("missing NCCL symbol ") + s; return p; };
            n.GetUniqueId = reinterpret_cast<decltype(n.GetUniqueId)>(sym("ncclGetUniqueId"));
            n.CommInitRank = reinterpret_cast<decltype(n.CommInitRank)>(sym("ncclCommInitRank"));
            n.CommDestroy = reinterpret_cast<decltype(n.CommDestroy)>(sym("ncclCommDestroy"));
            n.GroupStart = reinterpret_cast<decltype(n.GroupStart)>(sym("ncclGroupStart"));
            n.GroupEnd = reinterpret_cast<decltype(n.GroupEnd)>(sym("ncclGroupEnd"));
            n.Send = reinterpret_cast<decltype(n.Send)>(sym("ncclSend"));
            n.Recv = reinterpret_cast<decltype(n.Recv)>(sym("ncclRecv"));
            n.GetErrorString = reinterpret_cast<decltype(n.GetErrorString)>(sym("ncclGetErrorString"));
            n.ok = n.why.empty();
        });
        return n;
    }
};
constexpr int kNcclUint8 = 1;      // ncclDataType_t: ncclInt8 0, ncclUint8 1 (nccl.h)

}  // namespace

struct sr_ctx {
    int device = 0;
    cudaStream_t stream = nullptr, copy_stream = nullptr;
    std::string err;

    // mesh (device format: packed xyz, float4 rgb, int32 indices)
    DevBuf<float> pos;
    DevBuf<float4> col;
    DevBuf<int> idx;
    size_t nverts = 0, nidx = 0;
    unsigned long long first_bad = ~0ull;   // position of the first out-of-range index
    bool has_mesh = false;

    // per-frame working set
    DevBuf<float4> sv;                      // screen-space vertices (sx, sy, z, 0)
    DevBuf<TriRec> recs;                    // bin list (compact lists only)
    DevBuf<float4> ztri;                    // depth mode: corner depths of binned triangles
    DevBuf<TriRec> slab;                    // slab lists: ntiles x NCHUNK segments of slab_cap records
    DevBuf<unsigned> counts, cursor;        // (tile, chunk) counters / list cursors
    DevBuf<unsigned> pairs;                 // tile lists
    DevBuf<unsigned> vis32;                 // depth-off visibility words (prim+1), self-clearing
    DevBuf<unsigned long long> vis64;       // depth-less keys, self-clearing to SR_KEY_EMPTY
    DevBuf<unsigned long long> look;        // scan look-back array
    DevBuf<unsigned> ticket;
    unsigned ticket_base = 0, epoch = 0;
    DevBuf<unsigned> tile_done;             // frame stamp of tiles retired early (batched mode)
    DevBuf<unsigned> done_bits;             // the same as a bitmap (one bit per tile), all zero between frames
    unsigned frame_epoch = 0;
    // adaptive policy (depth off): index-descending batches pay off when most tile-list entries are never read
    bool batched_mode = false;
    unsigned long long single_batch_pairs = 0;
    std::vector<double> schedule;        // SR_SCHEDULE: forced batch shares (experiments); empty = two batches sized by first_frac
    double first_frac = 0.125;           // batched frames: share of the primitives (from the top) in the first batch
    int force_batches = -1;              // SR_BATCHES env: 0 = never, 1 = always, -1 = adaptive
    bool use_pdl = true;
    bool debug_sync = false;             // SR_DEBUG_SYNC=1: synchronise and report after every frame-kernel launch
    int max_spread = 0;                  // SR_SPREAD: largest spread_shift the binning k_setup may use on under-filled grids (0 = off)
    double spread_waves = 1.5;           // SR_SPREAD_WAVES: ... as long as the spread grid stays within this many waves
    int sm_count = 148;
    bool use_compact = false;            // SR_COMPACT=1: late batches of depth-off frames run the compacting k_setup (measured: slower)
    bool use_lbox = true;                // SR_LBOX=0: no box pre-pass in batched frames: the later batches transform every triangle themselves
                                         //   (the path frames with more than 255 tiles per axis take).  Measured on the 1M soup: 6,320 -> 6,657 frames/s
    DevBuf<unsigned> lbox;               // batched frames: clipped tile bounding box per primitive of the later batches (FrameParams::lbox)
    // speculative frames without the binning kernels: enabled after a frame in which no triangle was large; a frame
    // that does meet a large triangle only counts it and is re-run with the full pipeline (results identical)
    bool nobin_mode = false, call_nobin = false, allow_nobin = true;   // SR_NOBIN=0 disables
    // tile lists as fixed-capacity slabs (k_setup fills them directly; no k_scan / k_bin_fill).  A segment that
    // overflows makes the host double the capacity and re-run the frame; past kSlabLimit bytes the context falls
    // back to compact lists (scan + fill) for good.  SR_SLAB=0 forces compact lists.
    float z_lo = 0.0f, z_hi = 1.0f;      // depth mode: range of the triangles' depth bounds seen last frame (chunk thresholds)
    // two-pass depth frames (front pass, then the back pass culled against the front pass's per-tile farthest depth): a
    // launch policy like the others - chosen from the previous frame's counters, never changes the image
    bool depth_two_pass = false;
    int force_zpass = 0;                 // SR_ZPASS env: 0 = never (default), 1 = always, -1 = adaptive.  Default off: measured on the
                                         //   1M soup the two passes are bit-exact but SLOWER (1,914 vs 2,128 frames/s): k_setup 286 -> 202 us,
                                         //   k_tile 203 -> 303 us (the second tile pass costs more than the binning it saves)
    double zfrac = 0.125;                // SR_ZFRAC: share of the triangles (nearest corner first) the front pass takes
    float zsplit = 0.25f, zh_lo = 0.0f, zh_hi = 1.0f;   // split depth; range of the min-corner-depth histogram
    bool zsplit_valid = false;
    unsigned long long single_pass_pairs = 0;
    DevBuf<unsigned> tile_zmax;
    bool slab_mode = true;
    unsigned slab_cap = 32, call_slab_cap = 0;
    size_t last_ntiles = 0;
    DevBuf<ViewStatus> status;
    ViewStatus* h_status = nullptr;
    size_t h_status_cap = 0;

    // staging for the host-pointer API (double buffered)
    DevBuf<uint8_t> out_color[2];
    DevBuf<float> out_depth[2];
    cudaEvent_t ev_render[2] = {nullptr, nullptr}, ev_copy[2] = {nullptr, nullptr}, ev_clear[2] = {nullptr, nullptr};
    cudaEvent_t ev_begin = nullptr, ev_end = nullptr;

    // profiling
    bool profiling = false;
    cudaEvent_t prof_ev[MAX_PROF + 1] = {};
    const char* prof_name[MAX_PROF] = {};
    int prof_n = 0;
    float prof_ms[MAX_PROF] = {};

    // pending asynchronous views (sr_render_device .. sr_sync); several calls may accumulate before one sr_sync
    struct ViewJob { float m[16]; uint8_t* color; float* depth; };
    struct Pending {
        bool active = false;
        std::vector<ViewJob> jobs;
        int prim = 0, width = 0, height = 0, frame_height = 0, y_org = 0;   // height = rows rendered (the band's)
        uint32_t flags = 0;
        cudaStream_t stream = nullptr;
    } pending;

    // sr_set_band: render calls produce rows [band_y0, band_y1) of the frame only (band_y1 == 0: the whole frame)
    int band_y0 = 0, band_y1 = 0;

    sr_stats stats = {};
    uint64_t launches = 0;

    // host-side helpers of the host-pointer calls
    HostPool* pool = nullptr;            // SR_FILL_THREADS (default 4; 0 = constant depth plane crosses PCIe, pageable memory is left to the driver)
    const void* last_depth_ptr = nullptr;   // SR_FLAG_DEPTH_UNTOUCHED: what the previous host call left in the caller's depth image
    size_t last_depth_bytes = 0;
    bool last_depth_cleared = false;
    void* bounce[2] = {nullptr, nullptr};   // page-locked bounce buffers for pageable mesh arrays
    cudaEvent_t ev_bounce[2] = {nullptr, nullptr};
    DevBuf<uint8_t> mesh_stage;          // Swift-layout vertices + indices, converted on the device
    unsigned long long* h_bad = nullptr; // pinned: first out-of-range index position
    unsigned long long* d_bad = nullptr;
    uint64_t mesh_generation = 0;
    unsigned epoch0 = 0;                 // SR_EPOCH0 (tests): frame epoch a fresh tile_done array starts from

    // composite gather (sr_gather)
    Nccl::Comm comm = nullptr;
    int comm_rank = -1, comm_world = 0;
    cudaStream_t gather_stream = nullptr;
    cudaEvent_t ev_gather_in = nullptr, ev_gather_done[4] = {nullptr, nullptr, nullptr, nullptr};
    bool gather_slot_used[4] = {false, false, false, false};
};

namespace {

#define SR_CUDA(ctx, expr)                                                                             \
    do {                                                                                               \
        cudaError_t e__ = (expr);                                                                      \
        if (e__ != cudaSuccess) {                                                                      \
            char b__[512];                                                                             \
            snprintf(b__, sizeof b__, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__, __LINE__); \
            (ctx)->err = b__;                                                                          \
            return SR_ERR_CUDA;                                                                        \
        }                                                                                              \
    } while (0)

float depth_unord_host(unsigned o) {
    const unsigned u = (o & 0x80000000u) ? (o & 0x7fffffffu) : ~o;
    float f;
    std::memcpy(&f, &u, sizeof f);
    return f;
}

int fail(sr_ctx* c, int code, const char* msg) {
    if (c) c->err = msg;
    return code;
}

inline unsigned div_up(size_t a, size_t b) { return (unsigned)((a + b - 1) / b); }

void prof_mark(sr_ctx* c, cudaStream_t st, const char* name_of_next) {
    if (!c->profiling || c->prof_n >= MAX_PROF) return;
    cudaEventRecord(c->prof_ev[c->prof_n], st);
    c->prof_name[c->prof_n] = name_of_next;
    c->prof_n++;
}


// Every frame kernel is launched with programmatic stream serialization (PDL): its CTAs may be scheduled while the
// previous kernel drains, and block in griddepcontrol.wait (pdl_wait() at the top of each kernel) until that kernel
// has completed and flushed.  This hides launch latency between the many short kernels of a frame.  SR_PDL=0 disables.
template <class... KArgs, class... Args>
void launch_pdl(sr_ctx* c, void (*kernel)(KArgs...), unsigned grid, unsigned block, cudaStream_t st, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(block);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = c->use_pdl ? 1 : 0;
    cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
    if (c->debug_sync) {                  // SR_DEBUG_SYNC=1: find the launch that faults (launch ordinal of the call, grid, error)
        const cudaError_t e = cudaStreamSynchronize(st);
        fprintf(stderr, "[sr debug] launch %llu grid %u block %u: %s\n", (unsigned long long)c->launches, grid, block, cudaGetErrorString(e));
    }
}

int primitive_vertices(int prim) { return prim == SR_LINE ? 2 : 3; }   // R:179-188

// Every kernel of a frame asks for the SAME shared-memory carve-out of the unified L1/shared array.  Left to the
// driver's per-kernel choice, k_setup (24 KB per CTA, five CTAs) and k_tile (3 KB per CTA) get different splits: an SM
// cannot hold CTAs of two configurations at once, so the dependent kernel's CTAs cannot move in while the previous
// kernel drains (programmatic dependent launch is then void) and every kernel boundary costs a reconfiguration:
// 7-8 us between kernels measured with the %globaltimer timeline, against 0.8-1 us for equal configurations
// (tools/micro/pdl_gap.cu).  SR_CARVEOUT = percent of the array given to shared memory, -1 = driver default.
template <class K>
void set_carveout(K kernel, int pct) { cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, pct); }
void set_frame_carveouts(int pct) {
    if (pct < 0) return;
    set_carveout(k_transform, pct);
    set_carveout(k_setup<true, false, false>, pct);  set_carveout(k_setup<true, true, false>, pct);
    set_carveout(k_setup<false, false, false>, pct); set_carveout(k_setup<false, true, false>, pct);
    set_carveout(k_setup<false, true, true>, pct);   set_carveout(k_setup<false, true, true, true>, pct);
    set_carveout(k_setup<true, true, false, true>, pct); set_carveout(k_tile<true, true>, pct);
    set_carveout(k_setup<false, true, true, true, true>, pct); set_carveout(k_late_boxes, pct);
    set_carveout(k_scan, pct); set_carveout(k_bin_fill, pct);
    set_carveout(k_tile<true, false>, pct); set_carveout(k_tile<false, false>, pct); set_carveout(k_tile<false, true>, pct);
    set_carveout(k_resolve<true, 1>, pct); set_carveout(k_resolve<true, 4>, pct);
    set_carveout(k_resolve<false, 1>, pct); set_carveout(k_resolve<false, 4>, pct);
    set_carveout(k_points_scatter, pct); set_carveout(k_points_resolve, pct); set_carveout(k_clear_depth, pct);
}

// Buffers of the compact-list fallback: one record per triangle and the pair array (grown on demand by re-runs).
int ensure_compact(sr_ctx* c) {
    const size_t ntris = c->nidx / 3;
    SR_CUDA(c, c->recs.reserve(std::max<size_t>(ntris, 1)));
    if (!c->pairs.p) SR_CUDA(c, c->pairs.reserve(std::min<size_t>(std::max<size_t>(4 * ntris, 1u << 20), 64u << 20)));
    return SR_OK;
}

// Make sure the per-frame buffers fit this mesh and resolution.  Newly allocated self-clearing buffers are put
// into their cleared state here; afterwards the kernels keep them clean.
int ensure_frame_buffers(sr_ctx* c, int W, int H, bool depth_mode, size_t nviews, cudaStream_t st) {
    const size_t npx = (size_t)W * H;
    const size_t tiles = (size_t)div_up(W, TILE) * div_up(H, TILE);
    const size_t ncnt = tiles * NCHUNK;
    SR_CUDA(c, c->sv.reserve(std::max<size_t>(c->nverts, 1)));
    if (ncnt > c->counts.cap) {
        SR_CUDA(c, c->counts.reserve(ncnt));
        SR_CUDA(c, cudaMemsetAsync(c->counts.p, 0, ncnt * sizeof(unsigned), st));
    }
    SR_CUDA(c, c->cursor.reserve(ncnt));
    if (tiles > c->tile_done.cap) {
        SR_CUDA(c, c->tile_done.reserve(tiles));
        SR_CUDA(c, cudaMemsetAsync(c->tile_done.p, 0, tiles * sizeof(unsigned), st));
        c->frame_epoch = c->epoch0;
    }
    const size_t nbits_words = div_up(div_up(W, TILE), 32) * div_up(H, TILE);
    if (nbits_words > c->done_bits.cap) {
        SR_CUDA(c, c->done_bits.reserve(nbits_words));
        SR_CUDA(c, cudaMemsetAsync(c->done_bits.p, 0, nbits_words * sizeof(unsigned), st));
    }
    if (!depth_mode && c->batched_mode && c->use_lbox) SR_CUDA(c, c->lbox.reserve(std::max<size_t>(c->nidx / 3, 1)));
    const size_t nblocks = div_up(ncnt, SCAN_THREADS * SCAN_ITEMS);
    if (nblocks > c->look.cap) {
        SR_CUDA(c, c->look.reserve(nblocks));
        SR_CUDA(c, cudaMemsetAsync(c->look.p, 0, nblocks * sizeof(unsigned long long), st));
        c->epoch = 0;
    }
    if (!c->ticket.p) {
        SR_CUDA(c, c->ticket.reserve(1));
        SR_CUDA(c, cudaMemsetAsync(c->ticket.p, 0, sizeof(unsigned), st));
        c->ticket_base = 0;
    }
    if (c->slab_mode) {
        const size_t need = ncnt * c->slab_cap;
        if (need * sizeof(TriRec) > kSlabLimit) c->slab_mode = false;
        else SR_CUDA(c, c->slab.reserve(need));
    }
    if (!c->slab_mode) { int rc = ensure_compact(c); if (rc) return rc; }
    if (depth_mode) {
        SR_CUDA(c, c->tile_zmax.reserve(tiles));
        SR_CUDA(c, c->ztri.reserve(std::max<size_t>(c->nidx / 3, 1)));
        if (npx > c->vis64.cap) {
            SR_CUDA(c, c->vis64.reserve(npx));
            k_fill_u64<<<div_up(npx, 256), 256, 0, st>>>(c->vis64.p, npx, SR_KEY_EMPTY);
            SR_CUDA(c, cudaGetLastError());
        }
    }
    if (npx > c->vis32.cap) {   // also used by .vertices in either mode
        SR_CUDA(c, c->vis32.reserve(npx));
        SR_CUDA(c, cudaMemsetAsync(c->vis32.p, 0, npx * sizeof(unsigned), st));
    }
    SR_CUDA(c, c->status.reserve(nviews));
    if (nviews > c->h_status_cap) {
        if (c->h_status) cudaFreeHost(c->h_status);
        c->h_status = nullptr; c->h_status_cap = 0;
        SR_CUDA(c, cudaMallocHost((void**)&c->h_status, nviews * sizeof(ViewStatus)));
        c->h_status_cap = nviews;
    }
    return SR_OK;
}

// Rows of the frame a call renders: the band set by sr_set_band, else all H.
struct Band { int y0, rows; };
inline Band band_of(const sr_ctx* c, int H) { return c->band_y1 > 0 ? Band{c->band_y0, c->band_y1 - c->band_y0} : Band{0, H}; }

// Enqueue one render pass (R:204-230) on `st`.  Outputs are device pointers to images of `H` rows: rows
// [y_org, y_org + H) of the frame, whose full height `Hframe` only enters the viewport mapping (R:165-172).
int enqueue_view(sr_ctx* c, const float* transform, int prim, int W, int H, int Hframe, int y_org, uint8_t* color_dev, float* depth_dev,
                 uint32_t flags, cudaStream_t st, ViewStatus* d_status) {
    const bool depth_mode = (flags & SR_FLAG_DEPTH_LESS) != 0 && prim == SR_TRIANGLE;
    if (flags & SR_FLAG_NO_DEPTH_OUT) depth_dev = nullptr;
    const size_t npx = (size_t)W * H;
    Mat4 M;
    std::memcpy(M.m, transform, sizeof M.m);
    const int nverts = (int)c->nverts;
    unsigned* color32 = reinterpret_cast<unsigned*>(color_dev);
    c->prof_n = 0;

    // batched frames (depth off, heavy overdraw) transform corners inside k_setup, batch by batch
    const bool batched_frame = prim == SR_TRIANGLE && !depth_mode && c->batched_mode && (c->nidx / 3) >= (1u << 16);
    if (prim != SR_LINE && nverts > 0 && !batched_frame) {
        prof_mark(c, st, "k_transform");
        launch_pdl(c, k_transform, div_up(nverts, 256), 256, st, c->pos.p, nverts, M, (float)W, (float)Hframe, c->sv.p);
        c->launches++;
    }
    if (prim == SR_TRIANGLE) {
        const int ntris = (int)(c->nidx / 3);
        FrameParams fp;
        fp.width = W; fp.height = H; fp.y_org = y_org;
        fp.tiles_x = (int)div_up(W, TILE); fp.tiles_y = (int)div_up(H, TILE);
        fp.ntiles = fp.tiles_x * fp.tiles_y;
        fp.pair_capacity = (unsigned)std::min<size_t>(c->pairs.cap, 0xFFFFFFFFu);
        if (++c->frame_epoch == 0) {     // 2^32 frames: stamps of the previous cycle must not match the new epochs
            SR_CUDA(c, cudaMemsetAsync(c->tile_done.p, 0, c->tile_done.cap * sizeof(unsigned), st));
            c->frame_epoch = 1;
        }
        fp.epoch = c->frame_epoch;
        fp.batched = batched_frame ? 1 : 0;
        fp.done_bits = c->done_bits.p;
        fp.wpr = (int)div_up(fp.tiles_x, 32);
        fp.slab_cap = (c->slab_mode && (size_t)fp.ntiles * NCHUNK * c->slab_cap <= c->slab.cap) ? (int)c->slab_cap : 0;
        c->call_slab_cap = (unsigned)fp.slab_cap;
        c->last_ntiles = (size_t)fp.ntiles;
        for (int q = 0; q < NCHUNK - 1; ++q) fp.zthr[q] = c->z_lo + (c->z_hi - c->z_lo) * (float)(q + 1) / (float)NCHUNK;
        for (int q = 1; q < NCHUNK - 1; ++q) fp.zthr[q] = std::max(fp.zthr[q], fp.zthr[q - 1]);      // ascending, whatever the rounding
        fp.nobin = (c->nobin_mode && c->allow_nobin && !fp.batched) ? 1 : 0;
        if (fp.nobin) c->call_nobin = true;
        fp.tile_zmax = c->tile_zmax.p;
        // box pre-pass (FrameParams::lbox): tile coordinates must fit a byte, the buffer must hold one word per primitive
        fp.lbox = (fp.batched && c->use_lbox && !c->use_compact && c->max_spread == 0 && fp.tiles_x <= LBOX_MAX_TILES && fp.tiles_y <= LBOX_MAX_TILES &&
                   c->lbox.cap >= (size_t)ntris) ? c->lbox.p : nullptr;
        const bool two_pass = depth_mode && c->depth_two_pass && fp.slab_cap > 0 && !fp.nobin && ntris >= (1 << 16);
        const int ncnt = fp.ntiles * NCHUNK;
        BatchParams batches[MAX_BATCHES];
        int nb = 0;
        if (fp.batched) {
            // two index-descending batches: the top `first_frac` of the primitives (sized from how much of the tile
            // lists a single-batch frame consumed), then the rest, which mostly finds its tiles retired
            if (c->schedule.empty()) {
                const int first = std::min(ntris - 1, std::max(1, (int)((double)ntris * c->first_frac)));
                batches[nb++] = BatchParams{ntris - first, ntris, 0, 0, 0};
                batches[nb++] = BatchParams{0, ntris - first, 0, 1, 1};
            } else {
                // SR_SCHEDULE=f0,f1,...: shares of the primitives, from the top, of the first batches; the last batch takes the rest
                int hi = ntris;
                for (size_t q = 0; q < c->schedule.size() && nb < MAX_BATCHES - 1 && hi > 1; ++q) {
                    const int n = std::min(hi - 1, std::max(1, (int)((double)ntris * c->schedule[q])));
                    batches[nb] = BatchParams{hi - n, hi, 0, nb, 0};
                    hi -= n; nb++;
                }
                batches[nb] = BatchParams{0, hi, 0, nb, 1};
                nb++;
            }
        } else if (two_pass) {
            if (!c->zsplit_valid) { c->zsplit = c->zh_lo + 0.25f * (c->zh_hi - c->zh_lo); c->zsplit_valid = true; }
            const float scale = c->zh_hi > c->zh_lo ? (float)ZHIST_BINS / (c->zh_hi - c->zh_lo) : 0.0f;
            batches[nb++] = BatchParams{0, ntris, 0, 0, 0, 0, 1, c->zsplit, c->zh_lo, scale};
            batches[nb++] = BatchParams{0, ntris, 0, 1, 1, 0, 2, c->zsplit, c->zh_lo, scale};
        } else {
            batches[nb++] = BatchParams{0, ntris, 0, 0, 1};
        }
        ShadeIn in{c->idx.p, c->sv.p, c->col.p, c->ztri.p};
        for (int b = 0; b < nb; ++b) {
            BatchParams& bp = batches[b];
            const int n = bp.hi - bp.lo;
            while (n > 0 && ((n - 1) >> bp.chunk_shift) >= NCHUNK) bp.chunk_shift++;
            if (n > 0) {
                prof_mark(c, st, "k_setup");
                // later batches of a batched frame mostly meet retired tiles: the compacting variant (cheap test first,
                // full setup on the survivors with dense warps)
                const bool pre = fp.lbox != nullptr && b > 0;   // later batch of a frame with the box pre-pass: starts from its words
                const bool compact = (fp.batched && b > 0 && (c->use_compact || pre)) || two_pass;
                auto ks = depth_mode ? (fp.nobin ? k_setup<true, false, false> : (two_pass ? k_setup<true, true, false, true> : k_setup<true, true, false>))
                                     : (fp.nobin ? k_setup<false, false, false>
                                                 : (fp.batched ? (pre ? k_setup<false, true, true, true, true> : (compact ? k_setup<false, true, true, true> : k_setup<false, true, true>))
                                                               : k_setup<false, true, false>));
                XformIn xf{c->pos.p, M, (float)W, (float)Hframe};
                // an under-filled grid: fewer triangles per warp (BatchParams::spread_shift), while the grid stays within
                // spread_waves waves of 5 CTAs per SM.  SR_SPREAD=0 (default until measured) leaves it off.
                if (!compact && !fp.nobin && c->max_spread > 0) {
                    const size_t one_wave = (size_t)c->sm_count * 5 * SETUP_THREADS;
                    while (bp.spread_shift < c->max_spread && (double)((size_t)n << (bp.spread_shift + 1)) <= c->spread_waves * (double)one_wave) bp.spread_shift++;
                }
                const unsigned sgrid = div_up(compact ? (size_t)n : ((size_t)n << bp.spread_shift), compact ? SETUP_THREADS * (pre ? LBOX_PT : LATE_PT) : SETUP_THREADS);
                launch_pdl(c, ks, sgrid, SETUP_THREADS, st, c->idx.p, c->sv.p, xf, fp, bp, c->vis32.p, c->vis64.p, c->recs.p, c->counts.p, c->slab.p, c->ztri.p, c->tile_done.p, d_status);
                c->launches++;
                if (fp.lbox && b == 0 && bp.lo > 0) {        // the later batches' primitives [0, lo): boxes computed beside this k_setup
                    prof_mark(c, st, "k_late_boxes");
                    launch_pdl(c, k_late_boxes, div_up((size_t)bp.lo, 256 * PRE_PT), 256, st, c->idx.p, xf, fp, bp.lo, d_status);
                    c->launches++;
                }
            }
            if (!fp.nobin && !fp.slab_cap) {
                const unsigned nblocks = div_up(ncnt, SCAN_THREADS * SCAN_ITEMS);
                c->epoch++;
                prof_mark(c, st, "k_scan");
                launch_pdl(c, k_scan, nblocks, SCAN_THREADS, st, c->counts.p, c->cursor.p, ncnt, c->look.p, c->ticket.p, c->ticket_base, c->epoch, fp, bp.index, d_status);
                c->ticket_base += nblocks;
                c->launches++;
            }
            if (n > 0 && !fp.nobin && !fp.slab_cap) {
                prof_mark(c, st, "k_bin_fill");
                const unsigned grid = std::min<unsigned>(div_up(n, 256), 148 * 8);
                launch_pdl(c, k_bin_fill, grid, 256, st, c->recs.p, d_status, fp, bp, c->cursor.p, c->pairs.p, c->tile_done.p);
                c->launches++;
            }
            if (fp.nobin) {
                prof_mark(c, st, "k_resolve");
                // four pixels per thread once that still fills the GPU; its 16-byte loads and stores need the view's first
                // pixel on a 16-byte boundary in every plane (a view of a batch starts at a multiple of W*H elements; a
                // caller's device pointer may be a slice)
                const bool aligned16 = npx % 4 == 0 && (reinterpret_cast<uintptr_t>(color32) & 15) == 0 &&
                                       (!depth_dev || (reinterpret_cast<uintptr_t>(depth_dev) & 15) == 0);
                const bool wide = npx >= (4u << 20) && aligned16;
                auto kr = depth_mode ? (wide ? k_resolve<true, 4> : k_resolve<true, 1>) : (wide ? k_resolve<false, 4> : k_resolve<false, 1>);
                launch_pdl(c, kr, div_up(div_up(npx, wide ? 4 : 1), 256), 256, st, in, W, y_org, npx, c->vis32.p, c->vis64.p, color32, depth_dev);
                c->launches++;
                continue;
            }
            prof_mark(c, st, "k_tile");
            // late batches mostly find retired tiles: a persistent grid avoids thousands of empty CTA launches
            const unsigned tgrid = (fp.batched && b > 0) ? std::min<unsigned>(fp.ntiles, 148u * 6u)
                                 : ((two_pass && b > 0) ? std::min<unsigned>(fp.ntiles, 148u * 4u) : (unsigned)fp.ntiles);
            const TriRec* tile_recs = fp.slab_cap ? c->slab.p : c->recs.p;
            auto kt = depth_mode ? ((two_pass && b > 0) ? k_tile<true, true> : k_tile<true, false>)
                                 : ((fp.batched && b > 0) ? k_tile<false, true> : k_tile<false, false>);
            launch_pdl(c, kt, tgrid, 256, st, fp, bp, in, tile_recs, c->cursor.p, c->counts.p, c->pairs.p, c->vis32.p, c->vis64.p, c->tile_done.p, d_status, color32, depth_dev);
            c->launches++;
        }
    } else {
        if (prim == SR_VERTICES) {
            const int ncorners = (int)((c->nidx / 3) * 3);
            if (ncorners > 0) {
                prof_mark(c, st, "k_points_scatter");
                launch_pdl(c, k_points_scatter, div_up(ncorners, 256), 256, st, c->idx.p, c->sv.p, ncorners, W, H, y_org, c->vis32.p, d_status);
                c->launches++;
            }
        }
        prof_mark(c, st, "k_points_resolve");
        launch_pdl(c, k_points_resolve, div_up(npx, 256), 256, st, c->idx.p, c->col.p, npx, c->vis32.p, color32, depth_dev);
        c->launches++;
    }
    prof_mark(c, st, nullptr);
    SR_CUDA(c, cudaGetLastError());
    return SR_OK;
}

int check_args(sr_ctx* c, const float* transforms, size_t nviews, int prim, int W, int H, const void* color,
               const void* depth, uint32_t flags) {
    if (!c) return SR_ERR_ARG;
    if (!transforms || !color || nviews == 0) return fail(c, SR_ERR_ARG, "null transform/colour pointer or zero views");
    if (!depth && !(flags & SR_FLAG_NO_DEPTH_OUT)) return fail(c, SR_ERR_ARG, "depth pointer is null without SR_FLAG_NO_DEPTH_OUT");
    if (W <= 0 || H <= 0 || (size_t)W * H > (1u << 30)) return fail(c, SR_ERR_ARG, "bad image size");
    // what the packed fields of the binning kernels can represent: a bounding box's tile-column count lives in 15 bits
    // (StagedRect.w), tile numbers go through a float-estimated division that is exact below 2^23 (divmod_tiles)
    if ((W + TILE - 1) / TILE > 32767 || (size_t)((W + TILE - 1) / TILE) * ((H + TILE - 1) / TILE) >= (1u << 23))
        return fail(c, SR_ERR_ARG, "image too wide / too many 16x16 tiles for the binning kernels (width <= 524,272, tiles < 2^23)");
    if (prim == SR_VERTICES && c->nidx > 0x7FFFFFF0ull) return fail(c, SR_ERR_ARG, "too many corners for SR_VERTICES (32-bit corner numbers)");
    if (c->band_y1 > H) return fail(c, SR_ERR_ARG, "the band set by sr_set_band ends below the image");
    if (prim < SR_TRIANGLE || prim > SR_VERTICES) return fail(c, SR_ERR_ARG, "unknown primitive type");
    if (!c->has_mesh) return fail(c, SR_ERR_NOMESH, "sr_set_mesh has not been called");
    const size_t per = primitive_vertices(prim);
    const size_t used = (c->nidx / per) * per;                 // R:221: trailing partial primitive ignored
    if (c->first_bad < used) return fail(c, SR_ERR_INDEX, "index out of range (the reference traps at Renderer.swift:226)");
    return SR_OK;
}

// Reads the per-view status words back, grows the pair buffer and reports which views must be re-run.
int collect_status(sr_ctx* c, size_t nviews, cudaStream_t st, std::vector<size_t>* rerun) {
    SR_CUDA(c, cudaMemcpyAsync(c->h_status, c->status.p, nviews * sizeof(ViewStatus), cudaMemcpyDeviceToHost, st));
    SR_CUDA(c, cudaStreamSynchronize(st));
    unsigned need = 0;
    for (size_t i = 0; i < nviews; ++i) {
        const ViewStatus& s = c->h_status[i];
        unsigned most = 0;
        for (int b = 0; b < MAX_BATCHES; ++b) most = std::max(most, s.total_pairs[b]);
        // compact lists only: slab frames have no pair array (their k_tile still tallies list lengths into total_pairs)
        if (!c->call_slab_cap && !c->call_nobin && most > c->pairs.cap) { rerun->push_back(i); need = std::max(need, most); }
        else if (c->call_nobin && s.n_large[0] > 0) { rerun->push_back(i); c->stats.nobin_reruns++; }   // speculation failed
    }
    if (c->call_slab_cap) {
        unsigned seg = 0;
        for (size_t i = 0; i < nviews; ++i) {
            const unsigned m = c->h_status[i].seg_max;
            if (m > c->call_slab_cap) {
                if (std::find(rerun->begin(), rerun->end(), i) == rerun->end()) rerun->push_back(i);
                seg = std::max(seg, m);
            }
        }
        if (seg) {
            unsigned cap = c->slab_cap;
            while (cap < seg + seg / 4) cap *= 2;
            c->stats.bin_regrows++;
            if (c->last_ntiles * NCHUNK * (size_t)cap * sizeof(TriRec) > kSlabLimit) {
                c->slab_mode = false;
                c->slab.release();
                int rc = ensure_compact(c);
                if (rc) return rc;
            } else {
                c->slab_cap = cap;
                SR_CUDA(c, c->slab.reserve(c->last_ntiles * NCHUNK * (size_t)cap));
            }
        }
    }
    if (c->call_nobin && !rerun->empty()) c->nobin_mode = false;
    c->call_nobin = false;
    if (need > 0) {
        size_t want = (size_t)need + need / 4 + 1024;
        if (want > 0xFFFFFFF0ull) return fail(c, SR_ERR_CUDA, "tile lists exceed 2^32 entries");
        SR_CUDA(c, c->pairs.reserve(want));
        c->stats.bin_regrows++;
    }
    return SR_OK;
}

void accumulate_stats(sr_ctx* c, size_t nviews, int prim) {
    const size_t per = primitive_vertices(prim);
    for (size_t i = 0; i < nviews; ++i) {
        const ViewStatus& s = c->h_status[i];
        c->stats.primitives += c->nidx / per;
        c->stats.domain_skipped += s.domain_skipped;
        c->stats.small_tris += s.n_small;
        for (int b = 0; b < MAX_BATCHES; ++b) { c->stats.binned_tris += s.n_large[b]; c->stats.tile_pairs += s.total_pairs[b]; }
        c->stats.pairs_consumed += s.pairs_consumed;
        c->stats.depth_culled += s.pairs_culled;
        c->stats.batches_run += s.batches_run;
        c->stats.back_survivors += s.back_survivors;
        unsigned hist_total = 0;
        for (int q = 0; q < ZHIST_BINS; ++q) hist_total += s.zhist[q];
        if (hist_total) c->stats.depth_two_pass_frames++;
    }
    // adaptive policy for the next call (results are identical either way): batch when most list entries of a
    // single-batch frame were never read; go back when batching did not retire the frame early
    if (prim == SR_TRIANGLE && nviews > 0) {
        bool any_large = false;
        for (size_t i = 0; i < nviews; ++i)
            for (int b = 0; b < MAX_BATCHES; ++b) any_large |= c->h_status[i].n_large[b] != 0;
        c->nobin_mode = !any_large;
        {   // depth mode: next frame's front-to-back chunk thresholds span what this frame saw
            const ViewStatus& z = c->h_status[nviews - 1];
            if (z.zhi) {
                const float lo = depth_unord_host(~z.zlo_inv), hi = depth_unord_host(z.zhi);
                if (lo == lo && hi == hi && hi > lo) { c->z_lo = lo; c->z_hi = hi; }
            }
        }
        const ViewStatus& s = c->h_status[nviews - 1];
        {   // two-pass depth frames: on when most list entries of a single-pass frame were never used; the split depth is
            // the zfrac quantile of the nearest-corner depths, from the front pass's histogram
            unsigned long long hist_total = 0, pairs = 0;
            for (int q = 0; q < ZHIST_BINS; ++q) hist_total += s.zhist[q];
            for (int b = 0; b < MAX_BATCHES; ++b) pairs += s.total_pairs[b];
            const bool was_two_pass = hist_total != 0;
            if (was_two_pass) {
                unsigned long long cum = 0;
                const double target = c->zfrac * (double)hist_total;
                int bin = 0;
                while (bin < ZHIST_BINS - 1 && (double)(cum + s.zhist[bin]) < target) cum += s.zhist[bin++];
                // inside the bin: linear
                const double inbin = s.zhist[bin] ? std::min(1.0, std::max(0.0, (target - (double)cum) / (double)s.zhist[bin])) : 1.0;
                c->zsplit = c->zh_lo + (c->zh_hi - c->zh_lo) * (float)(((double)bin + inbin) / ZHIST_BINS);
                c->zsplit_valid = true;
                if (c->force_zpass < 0 && c->single_pass_pairs && pairs * 2 > c->single_pass_pairs) c->depth_two_pass = false;   // not saving list entries
            } else if (s.total_pairs[0] && (s.pairs_culled || s.pairs_consumed)) {          // a single-pass depth frame
                c->single_pass_pairs = s.total_pairs[0];
                if (c->force_zpass < 0) c->depth_two_pass = s.total_pairs[0] >= (1u << 18) && s.pairs_culled * 2 > s.total_pairs[0];
                if (!c->zsplit_valid) { c->zh_lo = c->z_lo; c->zh_hi = c->z_hi; }
            }
            if (c->force_zpass >= 0) c->depth_two_pass = c->force_zpass != 0;
            // the histogram range only ever grows (a wider range costs resolution, not correctness)
            if (c->z_lo < c->zh_lo) c->zh_lo = c->z_lo;
            if (c->z_hi > c->zh_hi) c->zh_hi = c->z_hi;
        }
        if (c->force_batches >= 0) c->batched_mode = c->force_batches != 0;
        else if (!c->batched_mode) {
            c->single_batch_pairs = s.total_pairs[0];
            c->batched_mode = s.total_pairs[0] >= (1u << 18) && s.pairs_consumed * 2 < s.total_pairs[0];
            // lists are consumed 16 records at a time, front first: 1.5x the consumed share covers the tiles that
            // needed more than the average (measured optimum on the 1M soup: 1/8 at a consumed share of 8 %)
            if (c->batched_mode)
                c->first_frac = std::min(0.5, std::max(1.0 / 64, 1.5 * (double)s.pairs_consumed / (double)s.total_pairs[0]));
        } else {
            unsigned long long sum = 0;
            for (int b = 0; b < MAX_BATCHES; ++b) sum += s.total_pairs[b];
            if (sum * 3 > c->single_batch_pairs) c->batched_mode = false;      // batching no longer saves binning work
        }
    }
}

// Debug build (-DSR_TIMELINE): one line per view with the overlapped per-launch timeline the kernels stamped into the
// status words (first CTA started, first CTA past the grid dependency, last CTA done), in µs from the frame's start.
void print_timeline(const sr_ctx* c, size_t nviews) {
#ifdef SR_TIMELINE
    for (size_t v = 0; v < nviews; ++v) {
        const ViewStatus& s = c->h_status[v];
        unsigned long long t0 = ~0ull;
        for (int k = 0; k < 2 * MAX_BATCHES; ++k) if (s.tl_launch_inv[k]) t0 = std::min(t0, ~s.tl_launch_inv[k]);
        fprintf(stderr, "view %zu:", v);
        for (int k = 0; k < 2 * MAX_BATCHES; ++k)
            if (s.tl_launch_inv[k])
                fprintf(stderr, "  [%s%d launch %.1f go %.1f end %.1f]", (k & 1) ? "tile" : "setup", k / 2, (~s.tl_launch_inv[k] - t0) * 1e-3,
                        (~s.tl_go_inv[k] - t0) * 1e-3, (s.tl_end[k] - t0) * 1e-3);
        fprintf(stderr, "\n");
        static const char* ph[7] = {"classified", "staged", "binned", "small done", "phase 1 done (compacting)", "phase 1: loads arrived", "phase 1: tests done"};
        for (int b = 0; b < 2; ++b)
            for (int k = 0; k < 7; ++k)
                if (s.ph_cnt[b][k])
                    fprintf(stderr, "  k_setup %s batch, warp reaches '%s': mean %.1f us, max %.1f us (%u warps)\n", b ? "later" : "first", ph[k],
                            (double)s.ph_sum[b][k] / s.ph_cnt[b][k] * 1e-3, (double)s.ph_max[b][k] * 1e-3, s.ph_cnt[b][k]);
    }
#else
    (void)c; (void)nviews;
#endif
}

void finish_profile(sr_ctx* c) {
    if (!c->profiling) return;
    for (int i = 0; i + 1 < c->prof_n; ++i) cudaEventElapsedTime(&c->prof_ms[i], c->prof_ev[i], c->prof_ev[i + 1]);
}

// Host -> device copy of caller memory.  Page-locked (sr_host_alloc / sr_host_register / cudaHostAlloc) memory is read by
// the copy engine directly.  Pageable memory goes through two page-locked bounce buffers, filled by the helper threads
// while the previous chunk is in flight (the driver's own staging of a pageable cudaMemcpyAsync is single-threaded).
int upload(sr_ctx* c, void* dst_dev, const void* src, size_t bytes, cudaStream_t st) {
    if (bytes == 0) return SR_OK;
    cudaPointerAttributes at = {};
    const bool pinned = cudaPointerGetAttributes(&at, src) == cudaSuccess && at.type == cudaMemoryTypeHost;
    cudaGetLastError();
    if (pinned || !c->pool || c->pool->threads() == 0 || bytes < (1u << 20)) {
        SR_CUDA(c, cudaMemcpyAsync(dst_dev, src, bytes, cudaMemcpyHostToDevice, st));
        return SR_OK;
    }
    constexpr size_t kChunk = 8u << 20;
    for (int b = 0; b < 2; ++b)
        if (!c->bounce[b]) SR_CUDA(c, cudaHostAlloc(&c->bounce[b], kChunk, cudaHostAllocPortable));
    int b = 0;
    for (size_t off = 0; off < bytes; off += kChunk, b ^= 1) {
        const size_t len = std::min(kChunk, bytes - off);
        SR_CUDA(c, cudaEventSynchronize(c->ev_bounce[b]));          // the copy that last read this buffer is done
        c->pool->start(HostPool::COPY, c->bounce[b], static_cast<const char*>(src) + off, len);
        c->pool->wait();
        SR_CUDA(c, cudaMemcpyAsync(static_cast<char*>(dst_dev) + off, c->bounce[b], len, cudaMemcpyHostToDevice, st));
        SR_CUDA(c, cudaEventRecord(c->ev_bounce[b], st));
    }
    return SR_OK;
}

// Completes the pending device-output views: status read-back, overflow re-runs, stats.  Synchronises `st`.
int finish_pending(sr_ctx* c) {
    auto& p = c->pending;
    const size_t nviews = p.jobs.size();
    cudaStream_t st = p.stream;
    for (int attempt = 0; attempt < 8; ++attempt) {
        std::vector<size_t> rerun;
        int rc = collect_status(c, nviews, st, &rerun);
        if (rc) return rc;
        if (rerun.empty()) break;
        if (attempt == 7) return fail(c, SR_ERR_CUDA, "speculative re-runs did not settle after 8 attempts (tile-list capacity)");
        for (size_t i : rerun) {
            SR_CUDA(c, cudaMemsetAsync(c->status.p + i, 0, sizeof(ViewStatus), st));
            rc = enqueue_view(c, p.jobs[i].m, p.prim, p.width, p.height, p.frame_height, p.y_org, p.jobs[i].color, p.jobs[i].depth, p.flags, st, c->status.p + i);
            if (rc) return rc;
        }
    }
    print_timeline(c, nviews);
    float ms = 0;
    cudaEventElapsedTime(&ms, c->ev_begin, c->ev_end);
    c->stats.gpu_ms = ms;
    finish_profile(c);
    accumulate_stats(c, nviews, p.prim);
    c->stats.gpu_launches = c->launches;
    p.jobs.clear();
    return c->stats.domain_skipped ? SR_ERR_DOMAIN : SR_OK;
}

}  // namespace

// =================================================================================================================
// C ABI
// =================================================================================================================
extern "C" {

int sr_version(void) { return 1000; }

int sr_create(sr_ctx** out, int device) {
    if (!out) return SR_ERR_ARG;
    *out = nullptr;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0 || device < 0 || device >= n) return SR_ERR_CUDA;   // no CPU fallback
    if (cudaSetDevice(device) != cudaSuccess) return SR_ERR_CUDA;
    sr_ctx* c = new sr_ctx();
    c->device = device;
    if (const char* e = getenv("SR_PDL")) c->use_pdl = atoi(e) != 0;
    if (const char* e = getenv("SR_DEBUG_SYNC")) c->debug_sync = atoi(e) != 0;
    {
        int pct = 57;                     // 146 KB of 256: five 24.5-KB k_setup CTAs (+1 KB each) fit
        if (const char* e = getenv("SR_CARVEOUT")) pct = atoi(e);
        set_frame_carveouts(pct);
        cudaGetLastError();
    }
    if (const char* e = getenv("SR_SPREAD")) c->max_spread = std::max(0, std::min(3, atoi(e)));
    if (const char* e = getenv("SR_SPREAD_WAVES")) c->spread_waves = std::max(0.5, atof(e));
    { int v = 0; if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, device) == cudaSuccess && v > 0) c->sm_count = v; }
    if (const char* e = getenv("SR_ZPASS")) { c->force_zpass = atoi(e); c->depth_two_pass = c->force_zpass > 0; }
    if (const char* e = getenv("SR_ZFRAC")) c->zfrac = std::min(0.9, std::max(0.005, atof(e)));
    if (const char* e = getenv("SR_COMPACT")) c->use_compact = atoi(e) != 0;
    if (const char* e = getenv("SR_LBOX")) c->use_lbox = atoi(e) != 0;
    if (const char* e = getenv("SR_SCHEDULE")) {
        for (const char* q = e; *q;) { char* end = nullptr; const double f = strtod(q, &end); if (end == q) break; if (f > 0.0 && f < 1.0) c->schedule.push_back(f); q = *end ? end + 1 : end; }
    }
    if (const char* e = getenv("SR_SLAB")) c->slab_mode = atoi(e) != 0;
    if (const char* e = getenv("SR_SLAB_CAP")) c->slab_cap = std::max(1, atoi(e));
    if (const char* e = getenv("SR_NOBIN")) c->allow_nobin = atoi(e) != 0;
    if (const char* e = getenv("SR_BATCHES")) { c->force_batches = atoi(e); c->batched_mode = c->force_batches > 0; }
    bool ok = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) == cudaSuccess &&
              cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking) == cudaSuccess;
    for (int i = 0; ok && i < 2; ++i)
        ok = cudaEventCreateWithFlags(&c->ev_render[i], cudaEventDisableTiming) == cudaSuccess &&
             cudaEventCreateWithFlags(&c->ev_copy[i], cudaEventDisableTiming) == cudaSuccess &&
             cudaEventCreateWithFlags(&c->ev_clear[i], cudaEventDisableTiming) == cudaSuccess;
    ok = ok && cudaEventCreate(&c->ev_begin) == cudaSuccess && cudaEventCreate(&c->ev_end) == cudaSuccess;
    for (int i = 0; ok && i <= MAX_PROF; ++i) ok = cudaEventCreate(&c->prof_ev[i]) == cudaSuccess;
    ok = ok && cudaStreamCreateWithFlags(&c->gather_stream, cudaStreamNonBlocking) == cudaSuccess &&
         cudaEventCreateWithFlags(&c->ev_gather_in, cudaEventDisableTiming) == cudaSuccess;
    for (int i = 0; ok && i < 4; ++i) ok = cudaEventCreateWithFlags(&c->ev_gather_done[i], cudaEventDisableTiming) == cudaSuccess;
    for (int i = 0; ok && i < 2; ++i) ok = cudaEventCreateWithFlags(&c->ev_bounce[i], cudaEventDisableTiming) == cudaSuccess;
    ok = ok && cudaMallocHost((void**)&c->h_bad, sizeof(unsigned long long)) == cudaSuccess &&
         cudaMalloc((void**)&c->d_bad, sizeof(unsigned long long)) == cudaSuccess;
    if (!ok) { sr_destroy(c); return SR_ERR_CUDA; }
    {
        int nthreads = 4;
        if (const char* e = getenv("SR_FILL_THREADS")) nthreads = std::max(0, std::min(64, atoi(e)));
        c->pool = new HostPool(nthreads);
    }
    if (const char* e = getenv("SR_EPOCH0")) c->epoch0 = (unsigned)strtoul(e, nullptr, 0);
    *out = c;
    return SR_OK;
}

void sr_destroy(sr_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    if (c->copy_stream) cudaStreamSynchronize(c->copy_stream);
    c->pos.release(); c->col.release(); c->idx.release(); c->sv.release(); c->recs.release(); c->slab.release(); c->ztri.release();
    c->tile_zmax.release(); c->lbox.release();
    c->counts.release(); c->cursor.release(); c->pairs.release(); c->vis32.release(); c->vis64.release();
    c->look.release(); c->ticket.release(); c->status.release(); c->tile_done.release(); c->done_bits.release();
    for (int i = 0; i < 2; ++i) {
        c->out_color[i].release(); c->out_depth[i].release();
        if (c->ev_render[i]) cudaEventDestroy(c->ev_render[i]);
        if (c->ev_copy[i]) cudaEventDestroy(c->ev_copy[i]);
        if (c->ev_clear[i]) cudaEventDestroy(c->ev_clear[i]);
    }
    if (c->ev_begin) cudaEventDestroy(c->ev_begin);
    if (c->ev_end) cudaEventDestroy(c->ev_end);
    for (int i = 0; i <= MAX_PROF; ++i) if (c->prof_ev[i]) cudaEventDestroy(c->prof_ev[i]);
    if (c->h_status) cudaFreeHost(c->h_status);
    delete c->pool;
    if (c->comm && Nccl::get().ok) Nccl::get().CommDestroy(c->comm);
    c->mesh_stage.release();
    for (int i = 0; i < 2; ++i) { if (c->bounce[i]) cudaFreeHost(c->bounce[i]); if (c->ev_bounce[i]) cudaEventDestroy(c->ev_bounce[i]); }
    for (int i = 0; i < 4; ++i) if (c->ev_gather_done[i]) cudaEventDestroy(c->ev_gather_done[i]);
    if (c->ev_gather_in) cudaEventDestroy(c->ev_gather_in);
    if (c->h_bad) cudaFreeHost(c->h_bad);
    if (c->d_bad) cudaFree(c->d_bad);
    if (c->gather_stream) { cudaStreamSynchronize(c->gather_stream); cudaStreamDestroy(c->gather_stream); }
    if (c->stream) cudaStreamDestroy(c->stream);
    if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
    delete c;
}

const char* sr_last_error(const sr_ctx* c) { return c ? c->err.c_str() : "null context"; }

int sr_set_mesh(sr_ctx* c, const sr_vertex* vertices, size_t nverts, const int64_t* indices, size_t nidx) {
    if (!c) return SR_ERR_ARG;
    if ((nverts && !vertices) || (nidx && !indices)) return fail(c, SR_ERR_ARG, "null mesh pointer");
    if (nverts > 0x7FFFFFF0ull || nidx > 3ull * 0x7FFFFFF0ull) return fail(c, SR_ERR_ARG, "mesh too large for 32-bit device indices");
    SR_CUDA(c, cudaSetDevice(c->device));
    if (c->pending.active) { int prc = sr_sync(c); if (prc == SR_ERR_CUDA) return prc; }
    SR_CUDA(c, cudaStreamSynchronize(c->stream));
    const bool same_shape = c->has_mesh && nverts == c->nverts && nidx == c->nidx;
    c->has_mesh = false;
    SR_CUDA(c, c->pos.reserve(std::max<size_t>(3 * nverts, 1)));
    SR_CUDA(c, c->col.reserve(std::max<size_t>(nverts, 1)));
    SR_CUDA(c, c->idx.reserve(std::max<size_t>(nidx, 1)));
    // The Swift-layout arrays (32-B vertices R:154-157, Int indices R:196) go into one persistent staging buffer and are
    // converted on the device.  Both copies and both conversions are enqueued back to back; the only host wait is the
    // one for the index check.  (The reference hands the arrays over on every call, A:163: sr_render_pass runs this
    // per frame, so nothing here allocates in the steady state.)
    const size_t vbytes = nverts * sizeof(sr_vertex), ibytes = nidx * sizeof(int64_t);
    const size_t voff = 0, ioff = (vbytes + 255) & ~(size_t)255;
    SR_CUDA(c, c->mesh_stage.reserve(std::max<size_t>(ioff + ibytes, 16)));
    *c->h_bad = ~0ull;
    SR_CUDA(c, cudaMemcpyAsync(c->d_bad, c->h_bad, sizeof(unsigned long long), cudaMemcpyHostToDevice, c->stream));
    if (nverts) {
        int rc = upload(c, c->mesh_stage.p + voff, vertices, vbytes, c->stream);
        if (rc) return rc;
        k_convert_vertices<<<div_up(nverts, 256), 256, 0, c->stream>>>((const float4*)(c->mesh_stage.p + voff), nverts, c->pos.p, c->col.p);
    }
    if (nidx) {
        int rc = upload(c, c->mesh_stage.p + ioff, indices, ibytes, c->stream);
        if (rc) return rc;
        k_convert_indices<<<div_up(nidx, 256), 256, 0, c->stream>>>((const long long*)(c->mesh_stage.p + ioff), nidx, nverts, c->idx.p, c->d_bad);
    }
    SR_CUDA(c, cudaGetLastError());
    SR_CUDA(c, cudaMemcpyAsync(c->h_bad, c->d_bad, sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
    SR_CUDA(c, cudaStreamSynchronize(c->stream));
    c->first_bad = *c->h_bad;
    c->nverts = nverts;
    c->nidx = nidx;
    c->has_mesh = true;
    c->mesh_generation++;
    if (!same_shape) {
        // the adaptive launch policies start over with a mesh of another size; a re-upload of the same shape (an
        // animated mesh, or the strict per-call hand-over of sr_render_pass) keeps them: they are hints that every frame
        // re-checks, results do not depend on them
        c->batched_mode = c->force_batches > 0;
        c->single_batch_pairs = 0;
        c->nobin_mode = false;
        c->z_lo = 0.0f; c->z_hi = 1.0f;
        c->depth_two_pass = c->force_zpass > 0;
        c->zsplit_valid = false;
        c->zh_lo = 0.0f; c->zh_hi = 1.0f;
        c->single_pass_pairs = 0;
    }
    return SR_OK;
}

uint64_t sr_mesh_generation(const sr_ctx* c) { return c ? c->mesh_generation : 0; }

int sr_host_alloc(sr_ctx* c, size_t bytes, void** out) {
    if (!c || !out || bytes == 0) return SR_ERR_ARG;
    *out = nullptr;
    SR_CUDA(c, cudaSetDevice(c->device));
    SR_CUDA(c, cudaHostAlloc(out, bytes, cudaHostAllocPortable));
    return SR_OK;
}

int sr_host_free(sr_ctx* c, void* ptr) {
    if (!c) return SR_ERR_ARG;
    if (!ptr) return SR_OK;
    SR_CUDA(c, cudaFreeHost(ptr));
    return SR_OK;
}

int sr_host_register(sr_ctx* c, void* ptr, size_t bytes) {
    if (!c || !ptr || bytes == 0) return SR_ERR_ARG;
    SR_CUDA(c, cudaSetDevice(c->device));
    SR_CUDA(c, cudaHostRegister(ptr, bytes, cudaHostRegisterPortable));
    return SR_OK;
}

int sr_host_fill_depth(sr_ctx* c, float* depth, size_t count) {
    if (!c || (!depth && count)) return SR_ERR_ARG;
    if (count == 0) return SR_OK;
    c->pool->start(HostPool::FILL_INF, depth, nullptr, count * sizeof(float));
    c->pool->wait();
    if (c->last_depth_ptr == depth) c->last_depth_cleared = false;     // not this library's render call: SR_FLAG_DEPTH_UNTOUCHED starts over
    return SR_OK;
}

int sr_host_unregister(sr_ctx* c, void* ptr) {
    if (!c || !ptr) return SR_ERR_ARG;
    SR_CUDA(c, cudaHostUnregister(ptr));
    return SR_OK;
}

// ---- composite gather over NCCL -------------------------------------------------------------------------------
int sr_comm_unique_id(void* id128_out) {
    if (!id128_out) return SR_ERR_ARG;
    Nccl& n = Nccl::get();
    if (!n.ok) return SR_ERR_COMM;
    Nccl::UniqueId id;
    if (n.GetUniqueId(&id) != 0) return SR_ERR_COMM;
    std::memcpy(id128_out, &id, sizeof id);
    return SR_OK;
}

int sr_comm_init(sr_ctx* c, const void* id128, int rank, int world) {
    if (!c || !id128 || world <= 0 || rank < 0 || rank >= world) return SR_ERR_ARG;
    Nccl& n = Nccl::get();
    if (!n.ok) return fail(c, SR_ERR_COMM, n.why.c_str());
    SR_CUDA(c, cudaSetDevice(c->device));
    if (c->comm) { n.CommDestroy(c->comm); c->comm = nullptr; }
    Nccl::UniqueId id;
    std::memcpy(&id, id128, sizeof id);
    const int rc = n.CommInitRank(&c->comm, world, id, rank);
    if (rc != 0) { c->comm = nullptr; c->err = std::string("ncclCommInitRank: ") + n.GetErrorString(rc); return SR_ERR_COMM; }
    c->comm_rank = rank;
    c->comm_world = world;
    return SR_OK;
}

int sr_comm_destroy(sr_ctx* c) {
    if (!c) return SR_ERR_ARG;
    if (c->comm) {
        cudaSetDevice(c->device);
        cudaStreamSynchronize(c->gather_stream);
        Nccl::get().CommDestroy(c->comm);
        c->comm = nullptr;
    }
    c->comm_rank = -1; c->comm_world = 0;
    return SR_OK;
}

int sr_gather(sr_ctx* c, const void* send_dev, void* recv_dev, const size_t* bytes_per_rank, int root, int slot, void* after_stream) {
    if (!c || !bytes_per_rank || slot < 0 || slot > 3) return SR_ERR_ARG;
    if (!c->comm) return fail(c, SR_ERR_COMM, "sr_gather before sr_comm_init");
    if (root < 0 || root >= c->comm_world) return fail(c, SR_ERR_ARG, "sr_gather: bad root");
    const size_t mine = bytes_per_rank[c->comm_rank];
    if ((mine && !send_dev) || (c->comm_rank == root && !recv_dev)) return fail(c, SR_ERR_ARG, "sr_gather: null buffer");
    Nccl& n = Nccl::get();
    SR_CUDA(c, cudaSetDevice(c->device));
    cudaStream_t after = after_stream ? (cudaStream_t)after_stream : c->stream;
    SR_CUDA(c, cudaEventRecord(c->ev_gather_in, after));
    SR_CUDA(c, cudaStreamWaitEvent(c->gather_stream, c->ev_gather_in, 0));
    int rc = 0;
    if (c->comm_rank == root) {
        // the root's own block is a device copy; the others arrive as one grouped batch of receives
        size_t off = 0;
        for (int r = 0; r < root; ++r) off += bytes_per_rank[r];
        if (mine && static_cast<char*>(recv_dev) + off != send_dev)
            SR_CUDA(c, cudaMemcpyAsync(static_cast<char*>(recv_dev) + off, send_dev, mine, cudaMemcpyDeviceToDevice, c->gather_stream));
        rc = n.GroupStart();
        off = 0;
        for (int r = 0; r < c->comm_world && rc == 0; ++r) {
            if (r != root && bytes_per_rank[r]) rc = n.Recv(static_cast<char*>(recv_dev) + off, bytes_per_rank[r], kNcclUint8, r, c->comm, c->gather_stream);
            off += bytes_per_rank[r];
        }
        const int rc2 = n.GroupEnd();
        if (rc == 0) rc = rc2;
    } else if (mine) {
        rc = n.Send(send_dev, mine, kNcclUint8, root, c->comm, c->gather_stream);
    }
    if (rc != 0) { c->err = std::string("sr_gather: ") + n.GetErrorString(rc); return SR_ERR_COMM; }
    SR_CUDA(c, cudaEventRecord(c->ev_gather_done[slot], c->gather_stream));
    c->gather_slot_used[slot] = true;
    return SR_OK;
}

int sr_gather_fence(sr_ctx* c, int slot, void* stream) {
    if (!c || slot < 0 || slot > 3) return SR_ERR_ARG;
    if (!c->gather_slot_used[slot]) return SR_OK;
    SR_CUDA(c, cudaSetDevice(c->device));
    SR_CUDA(c, cudaStreamWaitEvent(stream ? (cudaStream_t)stream : c->stream, c->ev_gather_done[slot], 0));
    return SR_OK;
}

int sr_gather_sync(sr_ctx* c) {
    if (!c) return SR_ERR_ARG;
    SR_CUDA(c, cudaSetDevice(c->device));
    SR_CUDA(c, cudaStreamSynchronize(c->gather_stream));
    return SR_OK;
}

int sr_render_device(sr_ctx* c, const float* transforms, size_t nviews, int prim, int W, int H, void* color_dev,
                     void* depth_dev, uint32_t flags, void* cuda_stream) {
    int rc = check_args(c, transforms, nviews, prim, W, H, color_dev, depth_dev, flags);
    if (rc) return rc;
    SR_CUDA(c, cudaSetDevice(c->device));
    cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : c->stream;
    const int Hframe = H;
    const Band band = band_of(c, H);
    H = band.rows;                                             // images hold the band's rows
    auto& p = c->pending;
    // calls with the same target format on the same stream accumulate (no host round trip between them)
    const size_t kStatusViews = 1024;
    if (p.active && (p.prim != prim || p.width != W || p.height != H || p.frame_height != Hframe || p.y_org != band.y0 || p.flags != flags || p.stream != st ||
                     p.jobs.size() + nviews > std::max(c->status.cap, kStatusViews))) {
        rc = sr_sync(c);
        if (rc == SR_ERR_CUDA) return rc;
    }
    const size_t first = p.active ? p.jobs.size() : 0;
    if (!p.active) {
        c->stats = sr_stats{};
        c->launches = 0;
        p.jobs.clear();
    }
    rc = ensure_frame_buffers(c, W, H, (flags & SR_FLAG_DEPTH_LESS) != 0, std::max(first + nviews, kStatusViews), st);
    if (rc) return rc;
    const size_t npx = (size_t)W * H;
    uint8_t* color = (uint8_t*)color_dev;
    float* depth = (flags & SR_FLAG_NO_DEPTH_OUT) ? nullptr : (float*)depth_dev;
    // one clear of every status slot when a group of accumulating calls starts (not one memset node per call)
    if (!p.active) {
        SR_CUDA(c, cudaMemsetAsync(c->status.p, 0, c->status.cap * sizeof(ViewStatus), st));
        SR_CUDA(c, cudaEventRecord(c->ev_begin, st));
    }
    p.active = true;
    p.prim = prim; p.width = W; p.height = H; p.frame_height = Hframe; p.y_org = band.y0; p.flags = flags; p.stream = st;
    for (size_t i = 0; i < nviews; ++i) {
        sr_ctx::ViewJob job;
        std::memcpy(job.m, transforms + 16 * i, sizeof job.m);
        job.color = color + 4 * npx * i;
        job.depth = depth ? depth + npx * i : nullptr;
        p.jobs.push_back(job);
        rc = enqueue_view(c, job.m, prim, W, H, Hframe, band.y0, job.color, job.depth, flags, st, c->status.p + first + i);
        if (rc) return rc;
    }
    SR_CUDA(c, cudaEventRecord(c->ev_end, st));
    return SR_OK;
}

int sr_sync(sr_ctx* c) {
    if (!c) return SR_ERR_ARG;
    if (!c->pending.active) return SR_OK;
    SR_CUDA(c, cudaSetDevice(c->device));
    c->pending.active = false;
    return finish_pending(c);
}

int sr_render_batch(sr_ctx* c, const float* transforms, size_t nviews, int prim, int W, int H, uint8_t* color,
                    float* depth, uint32_t flags) {
    int rc = check_args(c, transforms, nviews, prim, W, H, color, depth, flags);
    if (rc) return rc;
    SR_CUDA(c, cudaSetDevice(c->device));
    if (c->pending.active) { rc = sr_sync(c); if (rc == SR_ERR_CUDA) return rc; }
    c->stats = sr_stats{};
    c->launches = 0;
    cudaStream_t st = c->stream;
    const int Hframe = H;
    const Band band = band_of(c, H);
    H = band.rows;                                             // images hold the band's rows
    const bool want_depth = !(flags & SR_FLAG_NO_DEPTH_OUT);
    // Depth test off (or a primitive type that never tests depth): the pass only clears the depth image (R:206) and
    // never writes it again.  The clear then runs as its own kernel at the start of the frame and the cleared image
    // crosses PCIe while the frame renders; the tile kernel is told not to write depth.
    const bool early_depth = want_depth && !((flags & SR_FLAG_DEPTH_LESS) && prim == SR_TRIANGLE);
    const uint32_t view_flags = early_depth ? (flags | SR_FLAG_NO_DEPTH_OUT) : flags;
    const size_t npx = (size_t)W * H;
    // ... and a constant need not cross PCIe at all: the helper threads write +inf into the caller's image while the GPU
    // renders (streaming stores), or - on the caller's word that the image still holds the plane the previous call left
    // there (SR_FLAG_DEPTH_UNTOUCHED) - nothing is written.  SR_FILL_THREADS=0 keeps the PCIe copy.
    const size_t depth_bytes = nviews * npx * sizeof(float);
    const bool plane_kept = early_depth && (flags & SR_FLAG_DEPTH_UNTOUCHED) && c->last_depth_cleared &&
                            c->last_depth_ptr == depth && c->last_depth_bytes == depth_bytes;
    const bool host_fill = early_depth && !plane_kept && c->pool->threads() > 0;
    const bool pcie_plane = early_depth && !plane_kept && !host_fill;
    struct PoolWait { HostPool* p; ~PoolWait() { if (p) p->wait(); } } pool_wait{host_fill ? c->pool : nullptr};
    if (want_depth) { c->last_depth_ptr = depth; c->last_depth_bytes = depth_bytes; c->last_depth_cleared = false; }
    if (host_fill) c->pool->start(HostPool::FILL_INF, depth, nullptr, depth_bytes);
    rc = ensure_frame_buffers(c, W, H, (flags & SR_FLAG_DEPTH_LESS) != 0, nviews, st);
    if (rc) return rc;
    for (int b = 0; b < 2; ++b) {
        SR_CUDA(c, c->out_color[b].reserve(4 * npx));
        if (want_depth) SR_CUDA(c, c->out_depth[b].reserve(npx));
    }
    SR_CUDA(c, cudaMemsetAsync(c->status.p, 0, nviews * sizeof(ViewStatus), st));
    SR_CUDA(c, cudaEventRecord(c->ev_begin, st));
    // render view i into staging buffer i&1 while the copy stream drains the other one
    for (size_t i = 0; i < nviews; ++i) {
        const int b = (int)(i & 1);
        if (i >= 2) SR_CUDA(c, cudaStreamWaitEvent(st, c->ev_copy[b], 0));
        if (pcie_plane) {       // the render stream clears (it runs ahead of the copies), the copy stream only copies
            k_clear_depth<<<div_up(div_up(npx, 4), 256), 256, 0, st>>>(c->out_depth[b].p, npx);
            SR_CUDA(c, cudaGetLastError());
            c->launches++;
            SR_CUDA(c, cudaEventRecord(c->ev_clear[b], st));
            SR_CUDA(c, cudaStreamWaitEvent(c->copy_stream, c->ev_clear[b], 0));
            SR_CUDA(c, cudaMemcpyAsync(depth + npx * i, c->out_depth[b].p, npx * sizeof(float), cudaMemcpyDeviceToHost, c->copy_stream));
        }
        rc = enqueue_view(c, transforms + 16 * i, prim, W, H, Hframe, band.y0, c->out_color[b].p, (want_depth && !early_depth) ? c->out_depth[b].p : nullptr, view_flags, st, c->status.p + i);
        if (rc) return rc;
        SR_CUDA(c, cudaEventRecord(c->ev_render[b], st));
        SR_CUDA(c, cudaStreamWaitEvent(c->copy_stream, c->ev_render[b], 0));
        SR_CUDA(c, cudaMemcpyAsync(color + 4 * npx * i, c->out_color[b].p, 4 * npx, cudaMemcpyDeviceToHost, c->copy_stream));
        if (want_depth && !early_depth) SR_CUDA(c, cudaMemcpyAsync(depth + npx * i, c->out_depth[b].p, npx * sizeof(float), cudaMemcpyDeviceToHost, c->copy_stream));
        SR_CUDA(c, cudaEventRecord(c->ev_copy[b], c->copy_stream));
    }
    SR_CUDA(c, cudaEventRecord(c->ev_end, st));
    SR_CUDA(c, cudaStreamSynchronize(c->copy_stream));
    // status + overflow re-runs (rare: only while the tile-list buffer is still growing)
    for (int attempt = 0; attempt < 8; ++attempt) {
        std::vector<size_t> rerun;
        rc = collect_status(c, nviews, st, &rerun);
        if (rc) return rc;
        if (rerun.empty()) break;
        if (attempt == 7) return fail(c, SR_ERR_CUDA, "speculative re-runs did not settle after 8 attempts (tile-list capacity)");
        for (size_t i : rerun) {
            SR_CUDA(c, cudaMemsetAsync(c->status.p + i, 0, sizeof(ViewStatus), st));
            rc = enqueue_view(c, transforms + 16 * i, prim, W, H, Hframe, band.y0, c->out_color[0].p, (want_depth && !early_depth) ? c->out_depth[0].p : nullptr, view_flags, st, c->status.p + i);
            if (rc) return rc;
            SR_CUDA(c, cudaMemcpyAsync(color + 4 * npx * i, c->out_color[0].p, 4 * npx, cudaMemcpyDeviceToHost, st));
            if (want_depth && !early_depth) SR_CUDA(c, cudaMemcpyAsync(depth + npx * i, c->out_depth[0].p, npx * sizeof(float), cudaMemcpyDeviceToHost, st));
        }
    }
    print_timeline(c, nviews);
    float ms = 0;
    cudaEventElapsedTime(&ms, c->ev_begin, c->ev_end);
    c->stats.gpu_ms = ms;
    finish_profile(c);
    accumulate_stats(c, nviews, prim);
    c->stats.gpu_launches = c->launches;
    if (want_depth) c->last_depth_cleared = early_depth;
    return c->stats.domain_skipped ? SR_ERR_DOMAIN : SR_OK;
}

int sr_render(sr_ctx* c, const float* transform, int prim, int W, int H, uint8_t* color, float* depth, uint32_t flags) {
    return sr_render_batch(c, transform, 1, prim, W, H, color, depth, flags);
}

int sr_render_pass(sr_ctx* c, const sr_vertex* vertices, size_t nverts, const int64_t* indices, size_t nidx,
                   const float* transform, int prim, int W, int H, uint8_t* color, float* depth, uint32_t flags) {
    int rc = sr_set_mesh(c, vertices, nverts, indices, nidx);
    if (rc) return rc;
    return sr_render(c, transform, prim, W, H, color, depth, flags);
}

int sr_draw2d(sr_ctx* c, const sr_cmd2d* cmds, size_t n, int W, int H, uint8_t* color) {
    if (!c) return SR_ERR_ARG;
    if (!color || (n && !cmds) || W <= 0 || H <= 0 || (size_t)W * H > (1u << 30)) return fail(c, SR_ERR_ARG, "sr_draw2d: bad argument");
    auto sgn = [](int v) { return (v > 0) - (v < 0); };
    for (size_t k = 0; k < n; ++k) {                       // valid domain first: nothing is drawn if any command is outside it
        const sr_cmd2d& q = cmds[k];
        bool ok = q.kind >= SR_2D_RECT && q.kind <= SR_2D_TRIANGLE;
        if (!ok) return fail(c, SR_ERR_ARG, "sr_draw2d: unknown command");
        if (q.kind == SR_2D_RECT) {
            const long long x = q.p[0], y = q.p[1], w = q.p[2], h = q.p[3];
            ok = w != 0 && h != 0;
            if (ok) {
                const long long yb = y + w - sgn((int)w), xb = x + h - sgn((int)h);
                ok = std::min(y, yb) >= 0 && std::max(y, yb) < H && std::min(x, xb) >= 0 && std::max(x, xb) < W;
            }
        } else if (q.kind == SR_2D_CIRCLE) {
            const long long x = q.p[0], y = q.p[1], r = q.p[2];
            ok = r != 0;
            if (ok) {
                const long long a = -r, b = r - sgn((int)r);
                ok = y + std::min(a, b) >= 0 && y + std::max(a, b) < H && x + std::min(a, b) >= 0 && x + std::max(a, b) < W;
            }
        }
        if (!ok) return fail(c, SR_ERR_DOMAIN, "sr_draw2d: rect/circle footprint outside the image or zero extent "
                                               "(the reference writes through a raw pointer / traps, Renderer.swift:376-403)");
    }
    SR_CUDA(c, cudaSetDevice(c->device));
    if (c->pending.active) { int rc = sr_sync(c); if (rc == SR_ERR_CUDA) return rc; }
    cudaStream_t st = c->stream;
    const size_t npx = (size_t)W * H;
    SR_CUDA(c, c->out_color[0].reserve(4 * npx));
    unsigned* img = reinterpret_cast<unsigned*>(c->out_color[0].p);
    SR_CUDA(c, cudaMemcpyAsync(img, color, 4 * npx, cudaMemcpyHostToDevice, st));
    c->launches = 0;
    for (size_t k = 0; k < n; ++k) {
        const sr_cmd2d& q = cmds[k];
        unsigned col;
        std::memcpy(&col, q.color, 4);
        if (q.kind == SR_2D_RECT) {
            const size_t cnt = (size_t)std::abs((long long)q.p[2]) * (size_t)std::abs((long long)q.p[3]);
            k2d_rect<<<div_up(cnt, 256), 256, 0, st>>>(img, W, q.p[0], q.p[1], q.p[2], q.p[3], col);
        } else if (q.kind == SR_2D_CIRCLE) {
            const size_t side = 2 * (size_t)std::abs((long long)q.p[2]);
            k2d_circle<<<div_up(side * side, 256), 256, 0, st>>>(img, W, q.p[0], q.p[1], q.p[2]);
        } else if (q.kind == SR_2D_LINE) {
            k2d_line<<<1, 32, 0, st>>>(img, W, H, q.p[0], q.p[1], q.p[2], q.p[3], col);
        } else {
            const long long ylo = std::max<long long>(std::min({q.p[1], q.p[3], q.p[5]}), 0);
            const long long yhi = std::min<long long>(std::max({q.p[1], q.p[3], q.p[5]}), H - 1);
            if (yhi >= ylo) k2d_triangle<<<div_up((size_t)(yhi - ylo + 1), 128), 128, 0, st>>>(img, W, H, q.p[0], q.p[1], q.p[2], q.p[3], q.p[4], q.p[5]);
        }
        c->launches++;
    }
    SR_CUDA(c, cudaGetLastError());
    SR_CUDA(c, cudaMemcpyAsync(color, img, 4 * npx, cudaMemcpyDeviceToHost, st));
    SR_CUDA(c, cudaStreamSynchronize(st));
    c->stats = sr_stats{};
    c->stats.primitives = n;
    c->stats.gpu_launches = c->launches;
    return SR_OK;
}

int sr_set_band(sr_ctx* c, int y0, int y1) {
    if (!c) return SR_ERR_ARG;
    if (!(y0 == 0 && y1 == 0) && !(0 <= y0 && y0 < y1)) return fail(c, SR_ERR_ARG, "sr_set_band: need 0 <= y0 < y1, or (0, 0) for the whole frame");
    if (c->pending.active) { int rc = sr_sync(c); if (rc == SR_ERR_CUDA) return rc; }   // enqueued views keep the band they were given
    c->band_y0 = y0;
    c->band_y1 = y1;
    return SR_OK;
}

int sr_get_stats(sr_ctx* c, sr_stats* out) {
    if (!c || !out) return SR_ERR_ARG;
    *out = c->stats;
    return SR_OK;
}

int sr_set_profiling(sr_ctx* c, int enabled) {
    if (!c) return SR_ERR_ARG;
    c->profiling = enabled != 0;
    return SR_OK;
}

int sr_get_kernel_times(sr_ctx* c, const char** names, float* ms, int cap) {
    if (!c || !names || !ms) return 0;
    int n = 0;
    for (int i = 0; i + 1 < c->prof_n; ++i) {          // batches repeat kernels: sum by name
        int k = 0;
        while (k < n && std::strcmp(names[k], c->prof_name[i]) != 0) ++k;
        if (k == n) { if (n >= cap) continue; names[n] = c->prof_name[i]; ms[n] = 0.0f; ++n; }
        ms[k] += c->prof_ms[i];
    }
    return n;
}

}  // extern "C"

// =================================================================================================================
// Allocation-free entry point (include/sr_cuda.h: sr_points_no_alloc).  No counterpart in the reference.
// Nothing below calls cudaMalloc, cudaFree, cudaMallocHost or any allocating routine,
// and nothing above uses any of it: no existing code path changes.
// =================================================================================================================
namespace {

// The scratch: one visibility word per pixel (corner number + 1, 0 = background), SR_NOALLOC_MAX_PIXELS words = 4 MiB,
// held by every GPU that loads the library until the process exits.  In HOST code this name is a host-side stand-in,
// not a device address: kernels get the pointer cudaGetSymbolAddress returns (below), never `g_noalloc_vis` itself.
static __device__ __align__(256) unsigned g_noalloc_vis[SR_NOALLOC_MAX_PIXELS];

constexpr int kNoAllocMaxDevices = 64;
unsigned* g_noalloc_ptr[kNoAllocMaxDevices] = {};      // per device: the array's device address, resolved once
bool g_noalloc_ready[kNoAllocMaxDevices] = {};
std::mutex g_noalloc_mutex;
char g_noalloc_err[512] = "";

__global__ void __launch_bounds__(256) k_na_clear(unsigned* __restrict__ vis, unsigned n) {
    const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) vis[i] = 0u;
}

// R:295-302 per corner, straight from the Swift-layout arrays: Vertex.apply + convertedToScreen (R:159-172, the same
// device function the context path uses), Int(x), Int(y) truncating (R:298-299), bounds-checked store (R:30-36); the
// last corner in submission order wins = max(corner number + 1).
__global__ void __launch_bounds__(256) k_na_scatter(const float4* __restrict__ verts, unsigned long long nverts,
                                                    const long long* __restrict__ idx, unsigned ncorners, Mat4 M, int W, int H,
                                                    unsigned* __restrict__ vis) {
    const unsigned c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncorners) return;
    const long long i = idx[c];
    if (i < 0 || (unsigned long long)i >= nverts) return;                    // the reference traps (R:226): plots nothing here
    const float4 p = verts[2 * (size_t)i];
    const float3 s = transform_to_screen<false>(M, p.x, p.y, p.z, (float)W, (float)H);
    if (!(in_domain(s.x) && in_domain(s.y))) return;                         // the reference traps (R:298)
    const int x = __float2int_rz(s.x), y = __float2int_rz(s.y);
    if (x >= 0 && x < W && y >= 0 && y < H) atomicMax(vis + (size_t)y * W + x, c + 1u);
}

// Clears fused into the write-back (R:205-206), Pixel(float3: vertex.color) for the winner (R:300), and every word read
// is put back to zero, so the shared array is clean for the next call.
__global__ void __launch_bounds__(256) k_na_resolve(const float4* __restrict__ verts, const long long* __restrict__ idx, unsigned npx,
                                                    unsigned* __restrict__ vis, unsigned* __restrict__ color_out, float* __restrict__ depth_out) {
    const unsigned o = blockIdx.x * blockDim.x + threadIdx.x;
    if (o >= npx) return;
    const unsigned v = vis[o];
    unsigned bgra = 0u;
    if (v) {
        vis[o] = 0u;
        const float4 col = verts[2 * (size_t)idx[v - 1u] + 1];
        bgra = pack_bgra(col.x, col.y, col.z);
    }
    color_out[o] = bgra;
    if (depth_out) depth_out[o] = __int_as_float(0x7f800000);
}

int na_fail(int code, const char* what, cudaError_t e = cudaSuccess) {
    if (e != cudaSuccess) snprintf(g_noalloc_err, sizeof g_noalloc_err, "sr_points_no_alloc: %s: %s", what, cudaGetErrorString(e));
    else snprintf(g_noalloc_err, sizeof g_noalloc_err, "sr_points_no_alloc: %s", what);
    return code;
}

}  // namespace

extern "C" {

const char* sr_points_no_alloc_last_error(void) { return g_noalloc_err; }

int sr_points_no_alloc(int device, const sr_vertex* vertices_dev, size_t nverts, const int64_t* indices_dev, size_t nidx,
                       const float* transform, int W, int H, void* color_dev, void* depth_dev, void* cuda_stream) {
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);       // host-side state only: GPU work of overlapping calls still races (header)
    if (!transform || !color_dev || W <= 0 || H <= 0) return na_fail(SR_ERR_ARG, "null transform / colour pointer or non-positive size");
    if ((nverts && !vertices_dev) || (nidx && !indices_dev)) return na_fail(SR_ERR_ARG, "null mesh pointer");
    if ((unsigned long long)W * (unsigned long long)H > (unsigned long long)SR_NOALLOC_MAX_PIXELS) {
        snprintf(g_noalloc_err, sizeof g_noalloc_err, "sr_points_no_alloc: %d x %d = %llu pixels exceed the static scratch of %u pixels "
                 "(SR_NOALLOC_MAX_PIXELS); use sr_render* with a context for larger images", W, H, (unsigned long long)W * H, SR_NOALLOC_MAX_PIXELS);
        return SR_ERR_ARG;
    }
    if (nidx / 3 * 3 > 0xFFFFFFF0ull) return na_fail(SR_ERR_ARG, "too many corners (32-bit corner numbers)");
    if (device < 0 || device >= kNoAllocMaxDevices) return na_fail(SR_ERR_ARG, "device ordinal out of range");
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return na_fail(SR_ERR_CUDA, "cudaSetDevice", e);
    cudaStream_t st = (cudaStream_t)cuda_stream;
    const unsigned npx = (unsigned)W * (unsigned)H;
    if (!g_noalloc_ready[device]) {
        // once per device, with that device current: the array's DEVICE address (the symbol itself is a host stand-in)
        unsigned* ptr = nullptr;
        e = cudaGetSymbolAddress((void**)&ptr, g_noalloc_vis);
        if (e != cudaSuccess || !ptr) return na_fail(SR_ERR_CUDA, "cudaGetSymbolAddress(g_noalloc_vis)", e);
        g_noalloc_ptr[device] = ptr;
        k_na_clear<<<div_up(SR_NOALLOC_MAX_PIXELS, 256), 256, 0, st>>>(ptr, SR_NOALLOC_MAX_PIXELS);     // must start at zero
        if ((e = cudaGetLastError()) != cudaSuccess) return na_fail(SR_ERR_CUDA, "k_na_clear launch", e);
        g_noalloc_ready[device] = true;
    }
    unsigned* vis = g_noalloc_ptr[device];
    Mat4 M;
    std::memcpy(M.m, transform, sizeof M.m);
    const unsigned ncorners = (unsigned)(nidx / 3 * 3);                       // R:221: a trailing partial primitive is ignored
    if (ncorners)
        k_na_scatter<<<div_up(ncorners, 256), 256, 0, st>>>((const float4*)vertices_dev, (unsigned long long)nverts, (const long long*)indices_dev,
                                                            ncorners, M, W, H, vis);
    k_na_resolve<<<div_up(npx, 256), 256, 0, st>>>((const float4*)vertices_dev, (const long long*)indices_dev, npx, vis, (unsigned*)color_dev, (float*)depth_dev);
    if ((e = cudaGetLastError()) != cudaSuccess) return na_fail(SR_ERR_CUDA, "kernel launch", e);
    g_noalloc_err[0] = 0;
    return SR_OK;
}

}  // extern "C"
