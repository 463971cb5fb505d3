// sr_renderer.hpp — C++ host-side mirror of the reference's render interface over the C ABI (sr_cuda.h).
//
// The reference's host language is Swift; this image has no Swift toolchain, so the compiled-language host side is
// C++ (header-only).  Names and argument meaning follow Renderer.swift:
//   sr::Vertex         Renderer.swift:154-157      sr::PrimitiveType  Renderer.swift:174-189
//   sr::Pixel          Renderer.swift:44-49        sr::Image<T>       Renderer.swift:8-42
//   sr::RenderPass     Renderer.swift:191-200      sr::CudaRenderer::render(const RenderPass&)  Renderer.swift:204
// Error behaviour: where the reference traps (bad index R:226, NaN/inf corner R:251) render() throws
// std::runtime_error / returns the count of skipped primitives; it never falls back to the CPU.
#pragma once
#include <cstdint>
#include <initializer_list>
#include <stdexcept>
#include <string>
#include <vector>

#include "sr_cuda.h"

namespace sr {

struct float3 { float x, y, z, _pad; };                    // SIMD3<Float>: 16 bytes
struct Vertex { float3 xyz; float3 color; };               // 32 bytes
static_assert(sizeof(Vertex) == sizeof(sr_vertex), "Vertex must match Swift's layout");

struct Pixel { uint8_t b, g, r, a; };

enum class PrimitiveType : int { triangle = SR_TRIANGLE, line = SR_LINE, vertices = SR_VERTICES };
inline int verticesCount(PrimitiveType p) { return p == PrimitiveType::line ? 2 : 3; }

template <class T>
struct Image {                                             // non-owning view over caller memory
    T* pointer;
    int width, height, bytesPerRow;
};
using ColorImage = Image<Pixel>;
using DepthImage = Image<float>;

struct matrix_float4x4 {                                   // column-major, 64 bytes
    float columns[4][4] = {{1, 0, 0, 0}, {0, 1, 0, 0}, {0, 0, 1, 0}, {0, 0, 0, 1}};
};

struct RenderPass {
    ColorImage colorBuffer;
    DepthImage depthBuffer;
    std::vector<Vertex> vertices;
    std::vector<int64_t> indices;                          // Swift [Int]
    PrimitiveType primitiveType = PrimitiveType::triangle;
    matrix_float4x4 transform;
    // < 0 (default): the reference's contract - the arrays are handed over and uploaded on EVERY call (App.swift:163).
    // >= 0: "static mesh" opt-in - uploaded when this number or the array sizes differ from the upload before; the
    // caller bumps it after editing the arrays.  (Not a field of the reference's RenderPass.)
    long long meshVersion = -1;
};

class CudaRenderer {
public:
    explicit CudaRenderer(int device = 0) {
        int rc = sr_create(&ctx_, device);
        if (rc != SR_OK) throw std::runtime_error("sr_create failed (no CUDA device; there is no CPU fallback): " + std::to_string(rc));
    }
    ~CudaRenderer() { sr_destroy(ctx_); }
    CudaRenderer(const CudaRenderer&) = delete;
    CudaRenderer& operator=(const CudaRenderer&) = delete;

    bool depthTest = false;                                // true: the test commented out at Renderer.swift:257-261

    bool depthUntouched = false;                           // true: pass SR_FLAG_DEPTH_UNTOUCHED (see sr_cuda.h)

    // Explicit upload (sr_set_mesh), for passes with a meshVersion.
    void setMesh(const std::vector<Vertex>& v, const std::vector<int64_t>& i) {
        check(sr_set_mesh(ctx_, reinterpret_cast<const sr_vertex*>(v.data()), v.size(), i.data(), i.size()));
        mesh_version_ = -1;
    }

    // Page-locked host memory for images and mesh arrays (sr_host_alloc); free with hostFree.
    void* hostAlloc(size_t bytes) { void* p = nullptr; check(sr_host_alloc(ctx_, bytes, &p)); return p; }
    void hostFree(void* p) { check(sr_host_free(ctx_, p)); }

    // Screen-band split of one frame: later render() calls write only rows [y0, y1) of the pass's images (bit-identical
    // to those rows of the whole frame); setBand(0, 0) returns to whole frames.
    void setBand(int y0, int y1) { check(sr_set_band(ctx_, y0, y1)); band_y0_ = y1 > 0 ? y0 : 0; }

    // render(renderPass:) — synchronous; writes colorBuffer and depthBuffer in place.  Returns the number of
    // primitives skipped by the domain rule (0 for every input the reference itself can render).
    uint64_t render(const RenderPass& rp) {
        const size_t skip = static_cast<size_t>(band_y0_) * rp.colorBuffer.width;   // packed rows: a band is one block
        const uint32_t flags = (depthTest ? SR_FLAG_DEPTH_LESS : 0u) | (depthUntouched ? SR_FLAG_DEPTH_UNTOUCHED : 0u);
        uint8_t* color = reinterpret_cast<uint8_t*>(rp.colorBuffer.pointer + skip);
        float* depth = rp.depthBuffer.pointer + skip;
        int rc;
        if (rp.meshVersion < 0) {
            // no guessing from addresses or sizes: the arrays travel with the call, like the reference's
            mesh_version_ = -1;
            rc = sr_render_pass(ctx_, reinterpret_cast<const sr_vertex*>(rp.vertices.data()), rp.vertices.size(), rp.indices.data(),
                                rp.indices.size(), &rp.transform.columns[0][0], static_cast<int>(rp.primitiveType),
                                rp.colorBuffer.width, rp.colorBuffer.height, color, depth, flags);
        } else {
            if (rp.meshVersion != mesh_version_ || rp.vertices.size() != mesh_nv_ || rp.indices.size() != mesh_ni_ ||
                sr_mesh_generation(ctx_) != mesh_gen_) {
                setMesh(rp.vertices, rp.indices);
                mesh_version_ = rp.meshVersion; mesh_nv_ = rp.vertices.size(); mesh_ni_ = rp.indices.size();
                mesh_gen_ = sr_mesh_generation(ctx_);
            }
            rc = sr_render(ctx_, &rp.transform.columns[0][0], static_cast<int>(rp.primitiveType), rp.colorBuffer.width,
                           rp.colorBuffer.height, color, depth, flags);
        }
        sr_stats st{};
        sr_get_stats(ctx_, &st);
        if (rc != SR_OK && rc != SR_ERR_DOMAIN) check(rc);
        return st.domain_skipped;
    }

    // legacy 2-D immediate API (Renderer.swift:376-465): draw(rect:), draw(circle:), draw(line:), draw(triangle:with:in:)
    void drawRect(int x, int y, int w, int h, Pixel color, const ColorImage& image) { draw2d(SR_2D_RECT, {x, y, w, h, 0, 0}, color, image); }
    void drawCircle(int x, int y, int r, const ColorImage& image) { draw2d(SR_2D_CIRCLE, {x, y, r, 0, 0, 0}, Pixel{}, image); }
    void drawLine(int x0, int y0, int x1, int y1, Pixel color, const ColorImage& image) { draw2d(SR_2D_LINE, {x0, y0, x1, y1, 0, 0}, color, image); }
    void drawTriangle(int ax, int ay, int bx, int by, int cx, int cy, const ColorImage& image) { draw2d(SR_2D_TRIANGLE, {ax, ay, bx, by, cx, cy}, Pixel{}, image); }

    sr_stats stats() const { sr_stats st{}; sr_get_stats(ctx_, &st); return st; }
    sr_ctx* raw() const { return ctx_; }

private:
    void draw2d(int kind, std::initializer_list<int> p, Pixel color, const ColorImage& image) {
        sr_cmd2d cmd{};
        cmd.kind = kind;
        int k = 0;
        for (int v : p) cmd.p[k++] = v;
        cmd.color[0] = color.b; cmd.color[1] = color.g; cmd.color[2] = color.r; cmd.color[3] = color.a;
        check(sr_draw2d(ctx_, &cmd, 1, image.width, image.height, reinterpret_cast<uint8_t*>(image.pointer)));
    }
    void check(int rc) const {
        if (rc != SR_OK) throw std::runtime_error(std::string("libsr_cuda: ") + sr_last_error(ctx_) + " (" + std::to_string(rc) + ")");
    }
    sr_ctx* ctx_ = nullptr;
    long long mesh_version_ = -1;
    size_t mesh_nv_ = 0, mesh_ni_ = 0;
    uint64_t mesh_gen_ = 0;
    int band_y0_ = 0;
};

}  // namespace sr
