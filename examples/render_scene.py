#!/usr/bin/env python
"""Renders one of the BASELINE.json scenes through the CUDA path and writes a PNG — the headless stand-in for the
app's per-frame callback (App.swift:153-188: build RenderPass, call render(renderPass:), present).

    python examples/render_scene.py --scene cube --time 0.5 --out cube.png          (needs a B200)
    python examples/render_scene.py --scene soup --oracle --out soup_cpu.png        (CPU oracle, no GPU)
"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from software_renderer_b200 import scene as S  # noqa: E402

SCENES = {
    "cube": (lambda: S.cube(), 800, 600, lambda t: S.app_transform(t)),
    "icosphere": (lambda: S.icosphere(6), 1920, 1080, lambda t: S.app_transform(t)),
    "soup": (lambda: S.triangle_soup(100_000, 1920, 1080), 1920, 1080, lambda t: S.soup_transform(int(t * 60))),
    "cubesphere": (lambda: S.cube_sphere(289), 512, 512, lambda t: S.fibonacci_views(64)[int(t * 60) % 64]),
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scene", default="cube", choices=sorted(SCENES))
    ap.add_argument("--time", type=float, default=0.5)
    ap.add_argument("--depth-test", action="store_true")
    ap.add_argument("--oracle", action="store_true", help="render with the CPU oracle instead of the GPU")
    ap.add_argument("--out", default="frame.png")
    a = ap.parse_args()
    mesh, W, H, xf = SCENES[a.scene]
    v, i = mesh()
    m = xf(a.time)
    if a.oracle:
        from oracle import oracle as O
        rc, color, _, st = O.render(v, i, m, W, H, flags=O.DEPTH_LESS if a.depth_test else 0)
        print("oracle rc", rc, st)
    else:
        import software_renderer_b200 as srb
        with srb.CudaRenderer(0) as r:
            r.depth_test = a.depth_test
            rp = srb.RenderPass(srb.Image.color(W, H), srb.Image.depth(W, H), v, i, srb.PrimitiveType.triangle, m)
            r.render(rp)
            print({k: (int(x) if k != "gpu_ms" else round(float(x), 4)) for k, x in r.stats().items()})
            color = rp.color_buffer.array
    (S.write_ppm if a.out.endswith(".ppm") else S.write_png)(a.out, color)
    print("wrote", a.out, f"{W}x{H}, {int((np.asarray(color)[..., 3] > 0).sum())} covered pixels")


if __name__ == "__main__":
    main()
