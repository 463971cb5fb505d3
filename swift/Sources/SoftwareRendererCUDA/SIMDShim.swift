// SIMDShim.swift — the few Apple `simd` / CoreGraphics names Renderer.swift's public types use, for Linux.
// Only the TYPES cross the boundary; all arithmetic happens in libsr_cuda.so (and, for the oracle, in oracle/).
// UNVERIFIED: no Swift toolchain in this image.
#if !canImport(simd)
public typealias simd_float2 = SIMD2<Float>
public typealias simd_float3 = SIMD3<Float>
public typealias simd_float4 = SIMD4<Float>
public typealias vector_float2 = SIMD2<Float>
public typealias vector_float3 = SIMD3<Float>
public typealias vector_long2 = SIMD2<Int>
public typealias vector_long3 = SIMD3<Int>
public typealias simd_long2 = SIMD2<Int>

/// Column-major 4x4, the memory layout of Apple's matrix_float4x4 (64 bytes).
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
}
#else
import simd
#endif
