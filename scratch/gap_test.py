import sys, time, numpy as np, torch
sys.path.insert(0, '.')
import software_renderer_b200 as srb
from software_renderer_b200 import scene as S
W, H = 1920, 1080
v, i = S.triangle_soup(1_000_000, W, H)
r = srb.CudaRenderer(0); r.set_mesh(v, i)
n = 30
c = torch.empty((n, H, W, 4), dtype=torch.uint8, device='cuda'); d = torch.empty((n, H, W), dtype=torch.float32, device='cuda')
views = np.stack([S.soup_transform(k) for k in range(n)])
for rep in range(4):
    t0 = time.perf_counter()
    r.render_device(views, W, H, c.data_ptr(), d.data_ptr())
    t1 = time.perf_counter()
    s = r.stats()
    print(f"rep{rep}: {n} frames back-to-back gpu_ms/frame={s['gpu_ms']/n*1000:.1f} us  wall/frame={(t1-t0)/n*1e6:.1f} us launches/frame={s['gpu_launches']/n} batches_run={s['batches_run']/n}")
