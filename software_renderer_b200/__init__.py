"""software_renderer_b200 — B200 (sm_100a) implementation of the CPU rasterization path of
zhvrnkov/software-renderer (Renderer.render(renderPass:), Renderer.swift:204-230) behind a C ABI.

  csrc/           CUDA kernels + the C ABI (include/sr_cuda.h) -> libsr_cuda.so
  renderer.py     host-side mirror of the reference interface (RenderPass, PrimitiveType, CudaRenderer)
  scene.py        caller-side scene code: procedural meshes, baked colours, transform composition
"""
from .renderer import (Cmd2D, CudaRenderer, DomainError, Image, PrimitiveType, RenderPass, SrError, load_library,  # noqa: F401
                       ABI_SYMBOLS, LIB_PATH)
from . import scene  # noqa: F401
