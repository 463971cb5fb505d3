// swift-tools-version:5.7
// SwiftPM manifest of the Linux host side.  NOT BUILT OR TESTED IN THIS REPOSITORY'S IMAGE: there is no Swift toolchain
// here (SURVEY.md §0 row 7); every test goes through the same C ABI from Python/C++.  See INTEGRATION.md.
import PackageDescription

let package = Package(
    name: "SoftwareRendererCUDA",
    products: [.library(name: "SoftwareRendererCUDA", targets: ["SoftwareRendererCUDA"])],
    targets: [
        // C module map over include/sr_cuda.h; links libsr_cuda.so (software_renderer_b200/libsr_cuda.so)
        .systemLibrary(name: "CSRCuda", path: "Sources/CSRCuda"),
        // Linux stand-ins for Apple's `simd` and `Metal` modules (types + the three closed simd operations, DESIGN.md §2)
        .target(name: "simd", path: "Sources/simd"),
        .target(name: "Metal", path: "Sources/Metal"),
        // the reference's Renderer.swift, unchanged, as a second oracle: link the file in first (Sources/ReferenceOracle/README.md)
        .target(name: "ReferenceOracle", dependencies: ["simd", "Metal"], path: "Sources/ReferenceOracle", exclude: ["README.md"]),
        .target(name: "SoftwareRendererCUDA", dependencies: ["CSRCuda", "simd"],
                linkerSettings: [.unsafeFlags(["-L../software_renderer_b200", "-Xlinker", "-rpath", "-Xlinker", "../software_renderer_b200"])]),
    ]
)
