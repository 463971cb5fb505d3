/*
 * sr_cuda.h — C ABI of libsr_cuda.so, the B200 (sm_100a) implementation of the CPU rasterization path of
 * zhvrnkov/software-renderer.
 *
 * What it replaces ("R:" = renderer/Renderer.swift, "G:" = renderer/GpuRenderer.swift, "A:" = renderer/App.swift
 * in the reference tree):
 *   the seam is `func render(renderPass: RenderPass)` — Renderer R:204-230 (CPU) / GpuRenderer G:35-90 (Metal),
 *   called once per frame from A:185.  A Swift `CudaRenderer.render(renderPass:)` (swift/ in this repo) binds the
 *   functions below through a module map; data layouts are the reference's own, so the Swift
 *   arrays are passed without conversion:
 *     vertices : Swift `Vertex` R:154-157 = 2 x SIMD3<Float> = 32 bytes (x y z pad r g b pad)
 *     indices  : Swift `[Int]`  R:196     = int64
 *     transform: matrix_float4x4 R:199    = 16 floats, column-major
 *     colour   : `Pixel` R:44-49          = 4 bytes b,g,r,a; packed rows, y*width+x (A:351-360)
 *     depth    : Float per pixel (DepthImage R:6)
 *
 * Semantics are Renderer.swift's, bit for bit: clear to (0,0,0,0)/+inf (R:205-206), integer
 * scan-line coverage (R:271-283, R:467-494), "highest primitive index wins" visibility (the depth test is
 * commented out at R:257-261), barycentric colour (R:263-268).  SR_FLAG_DEPTH_LESS enables the commented-out
 * depth test as written there (strict <, so the earliest index wins ties).
 *
 * Error behaviour: the reference traps (R:226 index out of range, R:251/R:271/R:298 Int(Float) of NaN/inf/huge,
 * R:492 overflow).  This library never aborts: functions return 0 or a negative sr_status.  A primitive with a
 * corner whose screen x or y is not finite or has |.| >= 2^30 is skipped and counted; the frame is completed and
 * the call returns SR_ERR_DOMAIN.
 *
 * Threading: one sr_ctx per GPU; a context is single-threaded, like the reference renderer.
 * There is no CPU fallback: without a CUDA device sr_create fails with SR_ERR_CUDA.
 */
#ifndef SR_CUDA_H
#define SR_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct sr_ctx sr_ctx;

/* PrimitiveType, R:174-189 (verticesCount: triangle 3, line 2, vertices 3) */
typedef enum sr_primitive { SR_TRIANGLE = 0, SR_LINE = 1, SR_VERTICES = 2 } sr_primitive;

typedef enum sr_status {
    SR_OK = 0,
    SR_ERR_ARG = -1,     /* null pointer, non-positive size, unknown primitive type */
    SR_ERR_INDEX = -2,   /* an index is negative or >= nverts (reference: array trap, R:226) */
    SR_ERR_DOMAIN = -3,  /* >=1 primitive skipped: corner outside the valid domain (reference: trap, R:251) */
    SR_ERR_CUDA = -4,    /* CUDA runtime failure, or no device; see sr_last_error */
    SR_ERR_NOMESH = -5,  /* render called before sr_set_mesh */
    SR_ERR_COMM = -6     /* sr_gather: NCCL missing / not initialised / failed; see sr_last_error */
} sr_status;

enum {
    SR_FLAG_DEPTH_LESS = 1u << 0,   /* run the depth test of R:257-261 (off = reference as shipped) */
    SR_FLAG_NO_DEPTH_OUT = 1u << 1, /* do not produce / copy back the depth image (depth pointer may be NULL) */
    /* Host-pointer calls with the depth test off: the pass leaves the depth image at the clear value +inf (R:206; the
     * test that would write it is commented out, R:257-261).  With this flag the caller states that the depth image it
     * passes still holds what the previous call of this context wrote there (same pointer, same size, not touched
     * since): if that call left the cleared plane, the image is not written again.  Anything else (first call, another
     * pointer or size, a depth-tested call in between) is detected and the plane is written as usual. */
    SR_FLAG_DEPTH_UNTOUCHED = 1u << 2
};

/* Same bytes as Swift's Vertex (R:154-157). */
typedef struct sr_vertex {
    float x, y, z, _pad0;
    float r, g, b, _pad1;
} sr_vertex;

/* Per-context counters of the most recent render call (all views of a batch summed). */
typedef struct sr_stats {
    uint64_t primitives;      /* primitives submitted */
    uint64_t domain_skipped;  /* primitives (SR_VERTICES: corners) skipped by the domain rule */
    uint64_t small_tris;      /* triangles rasterized by the direct (per-triangle) path */
    uint64_t binned_tris;     /* triangles sent through tile binning */
    uint64_t tile_pairs;      /* (triangle, tile) pairs produced by binning */
    uint64_t bin_regrows;     /* times a frame was re-run because the pair buffer had to grow */
    uint64_t pairs_consumed;  /* pairs the tile kernel read before its exact early-out */
    uint64_t batches_run;     /* index-descending batches that still had open tiles (0 = single-batch frame) */
    uint64_t depth_culled;    /* depth mode: tile-list records rejected by their depth bound or never read */
    uint64_t nobin_reruns;    /* frames launched without the binning kernels that met a large triangle and were re-run */
    uint64_t depth_two_pass_frames; /* depth mode: views rendered as a front pass + a culled back pass */
    uint64_t back_survivors;  /* depth mode, two-pass views: triangles of the back pass that were not culled */
    uint64_t gpu_launches;    /* kernels launched by the call */
    float    gpu_ms;          /* device time of the call's kernels, CUDA events on the context's stream */
} sr_stats;

/* -- lifetime -------------------------------------------------------------------------------------------------- */
/* Replaces `Renderer()` / `GpuRenderer()` construction (A:148-149).  `device` is a CUDA ordinal. */
int sr_create(sr_ctx** out, int device);
void sr_destroy(sr_ctx* ctx);
const char* sr_last_error(const sr_ctx* ctx);
/* library / ABI version: major*1000 + minor */
int sr_version(void);

/* -- mesh -------------------------------------------------------------------------------------------------------
 * RenderPass.vertices / RenderPass.indices (R:195-196).  The reference hands both arrays over on every frame
 * (A:163); GpuRenderer caches its device buffers across frames (G:41-67).  Here the upload is explicit: call again
 * whenever the arrays change.  A trailing partial primitive is ignored, as R:221 does. */
int sr_set_mesh(sr_ctx* ctx, const sr_vertex* vertices, size_t nverts, const int64_t* indices, size_t nidx);
/* Generation counter of the uploaded mesh: incremented by every successful sr_set_mesh / sr_render_pass upload.  Host
 * mirrors that cache a mesh key on it together with a caller-bumped version (see CudaRenderer.setMesh). */
uint64_t sr_mesh_generation(const sr_ctx* ctx);

/* -- host memory -------------------------------------------------------------------------------------------------
 * The reference's caller owns the colour / depth images and the mesh arrays in ordinary memory (A:59-60, A:80-101,
 * A:163).  Every call below accepts such pointers.  Page-locked memory lets the GPU's copy engines write / read it
 * directly (about 3x the bandwidth of pageable memory on this box, profiles/): a host that wants that allocates its
 * images and arrays here, or registers the ones it already has.  Alignment requirements: none for host pointers;
 * device pointers (sr_render_device) need the natural alignment of their element (4 bytes). */
int sr_host_alloc(sr_ctx* ctx, size_t bytes, void** out);
int sr_host_free(sr_ctx* ctx, void* ptr);
int sr_host_register(sr_ctx* ctx, void* ptr, size_t bytes);
int sr_host_unregister(sr_ctx* ctx, void* ptr);
/* clear(image: depthBuffer, with: .infinity) (R:206, R:232-236) on caller memory, by the context's host threads with
 * streaming stores: `count` floats become +inf.  What the host-pointer calls do for the depth image of a pass that never
 * tests depth; exported so that a caller can clear an image itself (and bench.py can probe the box's ceiling for it). */
int sr_host_fill_depth(sr_ctx* ctx, float* depth, size_t count);

/* -- render ---------------------------------------------------------------------------------------------------- */
/* render(renderPass:) R:204-230 on the uploaded mesh: synchronous, writes host memory the caller owns
 * (A:80-101).  `color_bgra` = width*height*4 bytes, `depth` = width*height floats (may be NULL with
 * SR_FLAG_NO_DEPTH_OUT). */
int sr_render(sr_ctx* ctx, const float* transform_colmajor16, int primitive_type, int width, int height,
              uint8_t* color_bgra, float* depth, uint32_t flags);

/* One call = RenderPass as the reference passes it (vertices and indices travel with the call, A:163):
 * sr_set_mesh + sr_render. */
int sr_render_pass(sr_ctx* ctx, const sr_vertex* vertices, size_t nverts, const int64_t* indices, size_t nidx,
                   const float* transform_colmajor16, int primitive_type, int width, int height,
                   uint8_t* color_bgra, float* depth, uint32_t flags);

/* `nviews` independent render passes of the same mesh (one matrix each; the reference renders them one call at a
 * time).  Outputs are nviews consecutive images in host memory. */
int sr_render_batch(sr_ctx* ctx, const float* transforms_colmajor16, size_t nviews, int primitive_type,
                    int width, int height, uint8_t* color_bgra, float* depth, uint32_t flags);

/* Device-resident variant: outputs are DEVICE pointers (nviews consecutive images), work is enqueued on
 * `cuda_stream` (a cudaStream_t, NULL = the context's own stream) and the call returns without waiting.
 * `transforms_colmajor16` is host memory and is copied before return.  Follow with sr_sync.  Consecutive calls with the
 * same primitive type, size, flags and stream accumulate (no host round trip in between); one sr_sync completes them. */
int sr_render_device(sr_ctx* ctx, const float* transforms_colmajor16, size_t nviews, int primitive_type,
                     int width, int height, void* color_bgra_dev, void* depth_dev, uint32_t flags,
                     void* cuda_stream);

/* Screen-band split of ONE frame (tiles are independent, so a frame splits by rows).
 * After sr_set_band(ctx, y0, y1) every render call above produces only rows [y0, y1) of the `height`-row frame it
 * names: the colour / depth images it writes hold (y1 - y0) rows of `width` pixels (per view), bit-identical to
 * those rows of the whole frame (R:204-230 keeps no state across rows; the span rule R:467-494 depends on row
 * differences only, shading R:251-268 is evaluated in frame coordinates).  Ranks of a multi-GPU job render
 * disjoint bands of the same frame and gather them (multi_gpu.gather_bands).  (0, 0) returns to whole frames.
 * A band that ends below the image makes the render call fail with SR_ERR_ARG. */
int sr_set_band(sr_ctx* ctx, int y0, int y1);

/* Waits for the work enqueued by sr_render_device and returns its status (SR_OK / SR_ERR_DOMAIN / SR_ERR_CUDA). */
int sr_sync(sr_ctx* ctx);

/* -- composite: framebuffers of all ranks to one rank -------------------------------------------
 * Views shard over one process per GPU with no exchange while rendering (R:204-230 keeps no cross-frame state).  When
 * a composite is requested, each rank's finished framebuffers travel to `root` over NCCL (NVLink / NVSwitch) as
 * grouped ncclSend / ncclRecv (NCCL has no gather primitive).  libnccl.so.2 is opened at run time (the copy the
 * process already holds, e.g. PyTorch's, or the system one); without it these calls return SR_ERR_COMM.
 *   sr_comm_unique_id   rank `root` of the job creates the 128-byte NCCL id; the host side hands it to the other ranks
 *                       by whatever channel it has (torch.distributed / MPI / a file).
 *   sr_comm_init        every rank: joins the communicator (collective call).
 *   sr_gather           asynchronous: ordered after everything this context has enqueued so far on `after_stream` (NULL =
 *                       the context's own stream), it runs on the context's gather stream, so the gather of one chunk
 *                       of views overlaps the rendering of the next.  send_dev: this rank's block (device memory);
 *                       recv_dev (root only): where the blocks land, rank r's at byte offset sum(bytes_per_rank[0..r));
 *                       bytes_per_rank: `world` entries, identical on every rank.  `slot` (0..3) names the caller's
 *                       buffer for sr_gather_fence.
 *   sr_gather_fence     device-side: work enqueued on `stream` (NULL = own) after this call waits for the gather last
 *                       issued with this slot — call it before rendering into a buffer a gather may still be reading.
 *   sr_gather_sync      host-side: returns when every gather issued so far has completed. */
int sr_comm_unique_id(void* id128_out);
int sr_comm_init(sr_ctx* ctx, const void* id128, int rank, int world);
int sr_comm_destroy(sr_ctx* ctx);
int sr_gather(sr_ctx* ctx, const void* send_dev, void* recv_dev, const size_t* bytes_per_rank, int root, int slot,
              void* after_stream);
int sr_gather_fence(sr_ctx* ctx, int slot, void* stream);
int sr_gather_sync(sr_ctx* ctx);

/* -- allocation-free entry point ------------------------------------------------------------------------------------
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 *
 * One render pass of PrimitiveType.vertices (the arithmetic of R:295-302 after the clears of R:205-206: every corner
 * plots one pixel, the last corner in submission order wins, depth stays +inf) that allocates NOTHING: no sr_ctx, no
 * cudaMalloc / cudaFree / cudaMallocHost, no allocating library routine.  Its per-pixel visibility words live in a
 * file-scope `static __device__` array of SR_NOALLOC_MAX_PIXELS words (4 MiB) that every GPU which loads this library
 * holds until the process exits.  Inputs and outputs are the caller's DEVICE memory:
 *   vertices_dev   nverts Swift-layout vertices (sr_vertex, 32 bytes), 16-byte aligned
 *   indices_dev    nidx Swift `Int` indices (int64); corners used = (nidx / 3) * 3, as R:221; an index outside
 *                  [0, nverts) makes that corner plot nothing (the reference traps at R:226; checking the whole array
 *                  first would need a host round trip and scratch of its own)
 *   transform_colmajor16   HOST memory, 16 floats, passed to the kernel by value
 *   color_bgra_dev width*height*4 bytes; depth_dev width*height floats or NULL
 *   cuda_stream    a cudaStream_t, NULL = the default stream.  The call only enqueues: it returns without waiting.
 * Returns SR_ERR_ARG when width * height > SR_NOALLOC_MAX_PIXELS (or on null / non-positive arguments), SR_ERR_CUDA
 * on a CUDA failure, else SR_OK.  Corners outside the valid domain are skipped but NOT counted (no status word).
 * SHARED STATE: every call on a GPU uses the same array, which keeps its contents between calls; each call leaves it
 * all zero again (the resolve kernel re-arms every word it reads) and the first call on a device clears it.  Calls that
 * overlap on one GPU (two streams, two threads) RACE on it: serialise them.  sr_points_no_alloc_last_error() returns a
 * static string describing the last failure of this entry point in the calling process. */
#define SR_NOALLOC_MAX_PIXELS (1u << 20)
int sr_points_no_alloc(int device, const sr_vertex* vertices_dev, size_t nverts, const int64_t* indices_dev, size_t nidx,
                       const float* transform_colmajor16, int width, int height, void* color_bgra_dev, void* depth_dev,
                       void* cuda_stream);
const char* sr_points_no_alloc_last_error(void);

/* Counters of the last completed render call. */
int sr_get_stats(sr_ctx* ctx, sr_stats* out);

/* Per-kernel device times (ms) of the last single-view render, for bench.py's roofline line.
 * names: up to `cap` pointers to static strings; ms: same length.  Returns the number of entries. */
int sr_get_kernel_times(sr_ctx* ctx, const char** names, float* ms, int cap);
/* Enables per-kernel CUDA-event timing (adds events between kernels; off by default). */
int sr_set_profiling(sr_ctx* ctx, int enabled);

/* -- legacy 2-D immediate-mode API of Renderer.swift ------------------------------------ */
/* Replaces draw(rect:with:in:) R:376-383, draw(circle:with:in:) R:385-403, draw(line:with:in:) R:405-419 and
 * draw(triangle:with:in:) R:421-465.  The commands run in order on one caller-owned host BGRA image (in-out:
 * pixels no command touches keep their contents), each as a CUDA kernel on a device copy of the image.
 *   SR_2D_RECT     p = {x, y, w, h}: rows y ..< y+w step sign(w), columns x ..< x+h step sign(h) — the reference
 *                  swaps w and h (R:378-379) and so does this; unchecked stores of `color`.
 *   SR_2D_CIRCLE   p = {x, y, r}: dy, dx in -r ..< r, smooth-step edge, writes only the blue channel (R:401).
 *   SR_2D_LINE     p = {x0, y0, x1, y1}: float DDA with `.rounded()`, bounds-checked stores of `color`.
 *   SR_2D_TRIANGLE p = {ax, ay, bx, by, cx, cy}: integer corners, the span rule of R:467-494, colour = weights.
 * Error behaviour: the reference writes through a raw pointer for RECT and CIRCLE (out-of-image footprints corrupt
 * memory) and a zero w, h or r traps in stride(by: 0); such a command returns SR_ERR_DOMAIN and nothing is drawn. */
enum { SR_2D_RECT = 0, SR_2D_CIRCLE = 1, SR_2D_LINE = 2, SR_2D_TRIANGLE = 3 };
typedef struct sr_cmd2d {
    int32_t kind;
    int32_t p[6];
    uint8_t color[4]; /* b, g, r, a (Pixel, R:44-49) */
} sr_cmd2d;
int sr_draw2d(sr_ctx* ctx, const sr_cmd2d* cmds, size_t ncmds, int width, int height, uint8_t* color_bgra);

#ifdef __cplusplus
}
#endif
#endif
