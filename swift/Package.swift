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
        .target(name: "SoftwareRendererCUDA", dependencies: ["CSRCuda"],
                linkerSettings: [.unsafeFlags(["-L../software_renderer_b200", "-Xlinker", "-rpath", "-Xlinker", "../software_renderer_b200"])]),
    ]
)
