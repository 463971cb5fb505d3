#!/usr/bin/env python
"""Why does programmatic dependent launch not overlap kernels when the library runs inside the bench's Python process?
Runs a few frames with the -DSR_TIMELINE build (SR_LIB) under three setups and lets the library print its timeline:
  notorch   no torch in the process: device images from cudaMalloc through ctypes, the context's own stream
  own       torch imported and CUDA initialised by torch first, the context's own stream
  torch     torch's stream handed to sr_render_device (what bench.py does)
usage: SR_LIB=scratch/libsr_cuda_tl.so python tools/micro/py_pdl_probe.py notorch|own|torch"""
import ctypes
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
mode = sys.argv[1]
W, H = 1920, 1080
if mode != "notorch":
    import torch
    torch.cuda.set_device(0)
    torch.zeros(1, device="cuda")
import software_renderer_b200 as srb  # noqa: E402
from software_renderer_b200 import scene as S  # noqa: E402

v, i = S.triangle_soup(1_000_000, W, H)
r = srb.CudaRenderer(0)
r.depth_test = bool(os.environ.get("SR_PROBE_DEPTH"))      # SR_PROBE_DEPTH=1: the depth-tested mode's timeline
r.set_mesh(v, i)
if mode == "notorch":
    rt = ctypes.CDLL("libcudart.so")
    c, d = ctypes.c_void_p(), ctypes.c_void_p()
    assert rt.cudaMalloc(ctypes.byref(c), ctypes.c_size_t(4 * W * H)) == 0 and rt.cudaMalloc(ctypes.byref(d), ctypes.c_size_t(4 * W * H)) == 0
    cptr, dptr, sptr = c.value, d.value, None
else:
    color = torch.empty((H, W, 4), dtype=torch.uint8, device="cuda")
    depth = torch.empty((H, W), dtype=torch.float32, device="cuda")
    st = torch.cuda.Stream()
    cptr, dptr, sptr = color.data_ptr(), depth.data_ptr(), (st.cuda_stream if mode == "torch" else None)
mats = [r.colmajor(S.soup_transform(k)) for k in range(12)]
for k in range(6):
    r.render_device(mats[k], W, H, cptr, dptr, stream=sptr, sync=True, colmajor=True)
print(f"---- {mode}: timed frames", file=sys.stderr, flush=True)
for k in range(6, 12):
    r.render_device(mats[k], W, H, cptr, dptr, stream=sptr, sync=False, colmajor=True)
r.sync()
r.close()
