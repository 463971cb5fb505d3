// MetalShim.swift — a Linux module named `Metal` so that `import Metal` (Renderer.swift:3) resolves.  The reference's
// CPU path never touches the two Metal-typed optionals of Image (Renderer.swift:17-18); they only have to exist as types.
// UNVERIFIED: no Swift toolchain in this repository's image.
public protocol MTLTexture: AnyObject {}
public protocol MTLBuffer: AnyObject {}
