// SIMDShim.swift — a Linux module named `simd` with the Apple `simd` names Renderer.swift uses, so that the reference
// file compiles UNCHANGED (`import simd`, Renderer.swift:2) as a second oracle, and so that CudaRenderer's public types
// keep the reference's spelling.  UNVERIFIED: no Swift toolchain in this repository's image.
//
// Three operations live in Apple's closed library; their definitions here are the ones oracle/sr_oracle.cpp and
// csrc/sr_math.cuh share (DESIGN.md §2):
//   matrix * vector (Renderer.swift:160, :91)  column accumulation r = c0*x; r += c1*y; ...  Each product and sum is
//       rounded separately by default; building with -D SR_SIMD_FMA uses fused multiply-adds (Apple's simd_muladd on
//       arm64), the variant oracle/ and csrc/ build with -DSR_SIMD_FMA.
//   matrix_float2x2.inverse (Renderer.swift:99) adjugate / det, det = m00*m11 - m01*m10, every quotient a true division,
//       no singular early-out (an integer-degenerate triangle yields inf / NaN weights).
//   simd_clamp (Renderer.swift:119-122)         min(max(x, lo), hi) with fmax / fmin NaN handling (NaN -> lo).
@_exported import Foundation

public typealias simd_float2 = SIMD2<Float>
public typealias simd_float3 = SIMD3<Float>
public typealias simd_float4 = SIMD4<Float>
public typealias vector_float2 = SIMD2<Float>
public typealias vector_float3 = SIMD3<Float>
public typealias vector_float4 = SIMD4<Float>
public typealias vector_long2 = SIMD2<Int>
public typealias vector_long3 = SIMD3<Int>
public typealias simd_long2 = SIMD2<Int>
public typealias simd_long3 = SIMD3<Int>

@inline(__always) func sr_muladd(_ a: Float, _ b: Float, _ c: Float) -> Float {
#if SR_SIMD_FMA
    return c.addingProduct(a, b)          // one rounding (fma)
#else
    return a * b + c                      // Swift never contracts separate * and +
#endif
}

/// Column-major 2x2 (Renderer.swift:95-100, :141).
public struct matrix_float2x2 {
    public var columns: (simd_float2, simd_float2)
    public init(columns: (simd_float2, simd_float2)) { self.columns = columns }
    public init(diagonal d: simd_float2) { columns = (simd_float2(d.x, 0), simd_float2(0, d.y)) }
    public var inverse: matrix_float2x2 {
        let m00 = columns.0.x, m10 = columns.0.y, m01 = columns.1.x, m11 = columns.1.y
        let det = m00 * m11 - m01 * m10
        return matrix_float2x2(columns: (simd_float2(m11 / det, -m10 / det), simd_float2(-m01 / det, m00 / det)))
    }
    public static func * (m: matrix_float2x2, v: simd_float2) -> simd_float2 {
        let c0 = m.columns.0, c1 = m.columns.1
        return simd_float2(sr_muladd(c1.x, v.y, c0.x * v.x), sr_muladd(c1.y, v.y, c0.y * v.x))
    }
}

/// Column-major 4x4, the memory layout of Apple's matrix_float4x4 (64 bytes; Renderer.swift:199).
public struct matrix_float4x4 {
    public var columns: (simd_float4, simd_float4, simd_float4, simd_float4)
    public init(columns: (simd_float4, simd_float4, simd_float4, simd_float4)) { self.columns = columns }
    public init(diagonal d: simd_float4) {
        columns = (simd_float4(d.x, 0, 0, 0), simd_float4(0, d.y, 0, 0), simd_float4(0, 0, d.z, 0), simd_float4(0, 0, 0, d.w))
    }
    public init(rows r: [simd_float4]) {
        columns = (simd_float4(r[0].x, r[1].x, r[2].x, r[3].x), simd_float4(r[0].y, r[1].y, r[2].y, r[3].y),
                   simd_float4(r[0].z, r[1].z, r[2].z, r[3].z), simd_float4(r[0].w, r[1].w, r[2].w, r[3].w))
    }
    /// Renderer.swift:160 — r = c0*x; r = muladd(c1, y, r); r = muladd(c2, z, r); r = muladd(c3, w, r), per lane
    public static func * (m: matrix_float4x4, v: simd_float4) -> simd_float4 {
        var r = simd_float4(repeating: 0)
        for i in 0..<4 {
            var acc = m.columns.0[i] * v.x
            acc = sr_muladd(m.columns.1[i], v.y, acc)
            acc = sr_muladd(m.columns.2[i], v.z, acc)
            acc = sr_muladd(m.columns.3[i], v.w, acc)
            r[i] = acc
        }
        return r
    }
}

/// Renderer.swift:119-122, :396 — scalar clamp with fmax / fmin semantics: a NaN argument gives `lo`
public func simd_clamp(_ x: Float, _ lo: Float, _ hi: Float) -> Float { Float.minimum(Float.maximum(x, lo), hi) }
/// Renderer.swift:396 — t = clamp((x - e0) / (e1 - e0), 0, 1); t*t*(3 - 2t)
public func simd_smoothstep(_ e0: Float, _ e1: Float, _ x: Float) -> Float {
    let t = simd_clamp((x - e0) / (e1 - e0), 0, 1)
    return t * t * (3 - 2 * t)
}
/// Renderer.swift:390
public func length(_ v: simd_float2) -> Float { (v.x * v.x + v.y * v.y).squareRoot() }
