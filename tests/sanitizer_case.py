"""Small scene for `compute-sanitizer --tool racecheck|memcheck python tests/sanitizer_case.py` on the GPU box (not
collected by pytest): large, small and tiny triangles through k_setup / k_tile / k_resolve, depth off and on, whole
frame and two bands, each image checked against the oracle."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import software_renderer_b200 as srb  # noqa: E402
from oracle import oracle as O  # noqa: E402
from software_renderer_b200 import scene as S  # noqa: E402


def main():
    rng = np.random.default_rng(3)
    n = 160
    centre = rng.uniform(-1.0, 1.0, size=(n, 1, 3))
    ext = np.exp(rng.uniform(np.log(2e-3), np.log(1.2), size=(n, 1, 1)))
    xyz = (centre + rng.uniform(-1, 1, size=(n, 3, 3)) * ext).reshape(-1, 3)
    xyz[:, 2] = rng.uniform(0.0, 1.0, size=3 * n)
    v = S.pack_vertices(xyz.astype(np.float32), rng.uniform(0, 1, size=(3 * n, 3)).astype(np.float32))
    i = np.arange(3 * n, dtype=np.int64)
    m = np.eye(4, dtype=np.float32)
    W, H = 160, 112
    with srb.CudaRenderer(0) as r:
        r.set_mesh(v, i)
        for dt in (False, True):
            r.depth_test = dt
            rc, want_c, want_d, _ = O.render(v, i, m, W, H, flags=O.DEPTH_LESS if dt else 0)
            assert rc == O.OK
            for bands in ([(0, H)], [(0, 40), (40, H)]):
                parts = []
                for y0, y1 in bands:
                    r.set_band(y0, y1) if len(bands) > 1 else r.set_band()
                    parts.append(r.render_views(m[None], W, H))
                r.set_band()
                c = np.concatenate([p[0][0] for p in parts], axis=0)
                d = np.concatenate([p[1][0] for p in parts], axis=0)
                assert np.array_equal(c, want_c) and np.array_equal(d.view(np.uint32), want_d.view(np.uint32)), (dt, bands)
        small = S.icosphere(3)                                  # a frame without large triangles: k_resolve
        r.depth_test = False
        r.set_mesh(*small)
        for _ in range(2):
            c, d = r.render_views(S.app_transform(0.3)[None], W, H)
        rc, want_c, want_d, _ = O.render(small[0], small[1], S.app_transform(0.3), W, H)
        assert np.array_equal(c[0], want_c) and np.array_equal(d[0].view(np.uint32), want_d.view(np.uint32))
    print("sanitizer case ok")


if __name__ == "__main__":
    main()
