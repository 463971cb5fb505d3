// CudaRenderer.swift — drop-in for `Renderer` / `GpuRenderer`: same `render(renderPass:)` entry point over the same
// RenderPass / Vertex / PrimitiveType / Image types (Renderer.swift:8-49, 154-200 in the reference).  The caller picks
// the implementation by assignment (App.swift:154):   let renderer = cudaRenderer
// UNVERIFIED: this file cannot be compiled in this repository's image (no Swift toolchain).  The C ABI it binds is
// exercised by tests/ through ctypes.
import CSRCuda
import simd          // Apple's module on Darwin, Sources/simd on Linux

public typealias ColorImage = Image<Pixel>
public typealias DepthImage = Image<Float>

/// Renderer.swift:8-42 without the two Metal-only optionals (texture, buffer).
public final class Image<P> {
    public init(pointer: UnsafeMutablePointer<P>, width: Int, height: Int, bytesPerRow: Int) {
        self.pointer = pointer; self.width = width; self.height = height; self.bytesPerRow = bytesPerRow
    }
    public private(set) var pointer: UnsafeMutablePointer<P>
    public let width: Int, height: Int, bytesPerRow: Int
}

/// Renderer.swift:44-49
public struct Pixel { public var b: UInt8, g: UInt8, r: UInt8, a: UInt8 }

/// Renderer.swift:154-157 — two SIMD3<Float>: 32 bytes, the layout sr_vertex mirrors.
public struct Vertex {
    public let xyz: vector_float3
    public let color: vector_float3
    public init(xyz: vector_float3, color: vector_float3) { self.xyz = xyz; self.color = color }
}

/// Renderer.swift:174-189
public enum PrimitiveType {
    case triangle, line, vertices
    var raw: Int32 { switch self { case .triangle: return 0; case .line: return 1; case .vertices: return 2 } }
}

/// Renderer.swift:191-200
public struct RenderPass {
    public var colorBuffer: ColorImage
    public var depthBuffer: DepthImage
    public var vertices: [Vertex]
    public var indices: [Int]
    public var primitiveType: PrimitiveType = .triangle
    public var transform: matrix_float4x4 = .init(diagonal: .init(1, 1, 1, 1))
    /// nil (default) = the reference's contract: the arrays are handed over and uploaded on EVERY call (App.swift:163).
    /// A number = "static mesh" opt-in: uploaded when it (or an array count) differs from the upload before; the caller
    /// bumps it after editing `vertices` / `indices`.  Not a field of the reference's RenderPass.
    public var meshVersion: Int? = nil
    public init(colorBuffer: ColorImage, depthBuffer: DepthImage, vertices: [Vertex], indices: [Int]) {
        self.colorBuffer = colorBuffer; self.depthBuffer = depthBuffer; self.vertices = vertices; self.indices = indices
    }
}

/// Renderer.swift:51-70 — shapes of the legacy 2-D immediate API
public struct Rect { public var x: Int, y: Int, w: Int, h: Int; public init(x: Int, y: Int, w: Int, h: Int) { self.x = x; self.y = y; self.w = w; self.h = h } }
public struct Circle { public var x: Int, y: Int, r: Int; public init(x: Int, y: Int, r: Int) { self.x = x; self.y = y; self.r = r } }
public struct Line { public var x0: Int, y0: Int, x1: Int, y1: Int; public init(x0: Int, y0: Int, x1: Int, y1: Int) { self.x0 = x0; self.y0 = y0; self.x1 = x1; self.y1 = y1 } }
/// Renderer.swift:76-80
public struct Triangle { public var a: vector_long2, b: vector_long2, c: vector_long2; public init(a: vector_long2, b: vector_long2, c: vector_long2) { self.a = a; self.b = b; self.c = c } }

public enum CudaRendererError: Error { case create(Int32), render(Int32, String) }

public final class CudaRenderer {
    private var ctx: OpaquePointer?
    /// true = run the depth test that is commented out at Renderer.swift:257-261
    public var depthTest = false
    /// Set when the last pass skipped primitives outside the valid domain (the reference would have trapped).
    public private(set) var lastDomainSkips: UInt64 = 0
    /// true = pass SR_FLAG_DEPTH_UNTOUCHED: the depth image still holds what the previous call left there
    public var depthUntouched = false
    // static-mesh opt-in (RenderPass.meshVersion): what was uploaded last
    private var meshKey: (version: Int, nverts: Int, nidx: Int, generation: UInt64)? = nil

    public init(device: Int32 = 0) throws {
        precondition(MemoryLayout<Vertex>.stride == 32 && MemoryLayout<Int>.size == 8)
        let rc = sr_create(&ctx, device)
        guard rc == 0 else { throw CudaRendererError.create(rc) }   // no CPU fallback
    }
    deinit { sr_destroy(ctx) }

    /// Screen-band split of one frame: later render(renderPass:) calls write only rows y0 ..< y1 of the pass's images,
    /// bit-identical to those rows of the whole frame; setBand(0, 0) returns to whole frames.
    private var bandStart = 0
    public func setBand(_ y0: Int, _ y1: Int) {
        let rc = sr_set_band(ctx, Int32(y0), Int32(y1))
        precondition(rc == 0, String(cString: sr_last_error(ctx)))
        bandStart = y1 > 0 ? y0 : 0
    }

    /// Same signature and contract as Renderer.render(renderPass:) (Renderer.swift:204): synchronous, writes
    /// colorBuffer and depthBuffer in place.
    public func render(renderPass: RenderPass) {
        let w = Int32(renderPass.colorBuffer.width), h = Int32(renderPass.colorBuffer.height)
        var m = renderPass.transform
        let flags: UInt32 = (depthTest ? UInt32(SR_FLAG_DEPTH_LESS) : 0) | (depthUntouched ? UInt32(SR_FLAG_DEPTH_UNTOUCHED) : 0)
        let skip = bandStart * renderPass.colorBuffer.width          // packed rows: a band is one block of the image
        let color = UnsafeMutableRawPointer(renderPass.colorBuffer.pointer + skip).assumingMemoryBound(to: UInt8.self)
        let depth = renderPass.depthBuffer.pointer + skip
        var rc: Int32 = 0
        if let version = renderPass.meshVersion {
            // static mesh: upload on a version / size change only (no hashing, no guessing from addresses)
            let key = (version, renderPass.vertices.count, renderPass.indices.count, sr_mesh_generation(ctx))
            if meshKey == nil || meshKey! != key {
                setMesh(vertices: renderPass.vertices, indices: renderPass.indices)
                meshKey = (version, renderPass.vertices.count, renderPass.indices.count, sr_mesh_generation(ctx))
            }
            rc = withUnsafeBytes(of: &m) { mb in
                sr_render(ctx, mb.baseAddress!.assumingMemoryBound(to: Float.self), renderPass.primitiveType.raw, w, h, color, depth, flags)
            }
        } else {
            // the reference's contract: vertices and indices travel with every call (App.swift:163) -> sr_render_pass
            meshKey = nil
            rc = renderPass.vertices.withUnsafeBytes { vb in
                renderPass.indices.withUnsafeBufferPointer { ib in
                    withUnsafeBytes(of: &m) { mb in
                        sr_render_pass(ctx, vb.baseAddress?.assumingMemoryBound(to: sr_vertex.self), renderPass.vertices.count,
                                       UnsafeRawPointer(ib.baseAddress)?.assumingMemoryBound(to: Int64.self), renderPass.indices.count,
                                       mb.baseAddress!.assumingMemoryBound(to: Float.self), renderPass.primitiveType.raw, w, h, color, depth, flags)
                    }
                }
            }
        }
        var st = sr_stats()
        sr_get_stats(ctx, &st)
        lastDomainSkips = st.domain_skipped
        precondition(rc == 0 || rc == SR_ERR_DOMAIN.rawValue, String(cString: sr_last_error(ctx)))   // Renderer.swift:226 traps on a bad index
    }

    /// Explicit upload (sr_set_mesh) for passes that carry a `meshVersion`.
    public func setMesh(vertices: [Vertex], indices: [Int]) {
        let rc = vertices.withUnsafeBytes { vb in
            indices.withUnsafeBufferPointer { ib in
                sr_set_mesh(ctx, vb.baseAddress?.assumingMemoryBound(to: sr_vertex.self), vertices.count,
                            UnsafeRawPointer(ib.baseAddress)?.assumingMemoryBound(to: Int64.self), indices.count)
            }
        }
        precondition(rc == 0, String(cString: sr_last_error(ctx)))
        meshKey = nil
    }

    /// Page-locked memory for the colour / depth images (sr_host_alloc): the GPU's copy engines write it directly.
    public func hostAlloc<T>(_ type: T.Type, count: Int) -> UnsafeMutablePointer<T> {
        var p: UnsafeMutableRawPointer? = nil
        let rc = sr_host_alloc(ctx, count * MemoryLayout<T>.stride, &p)
        precondition(rc == 0 && p != nil, String(cString: sr_last_error(ctx)))
        return p!.bindMemory(to: T.self, capacity: count)
    }
    public func hostFree<T>(_ p: UnsafeMutablePointer<T>) { _ = sr_host_free(ctx, UnsafeMutableRawPointer(p)) }

    // MARK: legacy 2-D immediate API (Renderer.swift:376-465) — same signatures, one sr_draw2d command each

    private func draw2d(_ kind: Int32, _ p: [Int], _ color: Pixel, in image: ColorImage) {
        var cmd = sr_cmd2d()
        cmd.kind = kind
        withUnsafeMutableBytes(of: &cmd.p) { b in for (k, v) in p.enumerated() { b.storeBytes(of: Int32(truncatingIfNeeded: v), toByteOffset: 4 * k, as: Int32.self) } }
        cmd.color = (color.b, color.g, color.r, color.a)
        let rc = sr_draw2d(ctx, &cmd, 1, Int32(image.width), Int32(image.height),
                           UnsafeMutableRawPointer(image.pointer).assumingMemoryBound(to: UInt8.self))
        precondition(rc == 0, String(cString: sr_last_error(ctx)))   // the reference corrupts memory / traps here
    }
    public func draw(rect: Rect, with color: Pixel, in image: ColorImage) {
        draw2d(Int32(SR_2D_RECT.rawValue), [rect.x, rect.y, rect.w, rect.h], color, in: image)
    }
    public func draw(circle: Circle, with color: Pixel, in image: ColorImage) {
        draw2d(Int32(SR_2D_CIRCLE.rawValue), [circle.x, circle.y, circle.r], color, in: image)
    }
    public func draw(line: Line, with color: Pixel, in image: ColorImage) {
        draw2d(Int32(SR_2D_LINE.rawValue), [line.x0, line.y0, line.x1, line.y1], color, in: image)
    }
    public func draw(triangle: Triangle, with color: Pixel, in image: ColorImage) {
        draw2d(Int32(SR_2D_TRIANGLE.rawValue), [triangle.a.x, triangle.a.y, triangle.b.x, triangle.b.y, triangle.c.x, triangle.c.y], color, in: image)
    }
}
