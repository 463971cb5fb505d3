#!/usr/bin/env python
"""Multi-GPU check run by hand on a box with >= 2 GPUs (not collected by pytest: it needs torchrun):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 \
        tests/nccl_composite_check.py

Each rank renders its block of the view list on its own GPU (no collective on the data path), the finished
framebuffers are gathered to rank 0 over NCCL, and rank 0 compares every view bit for bit with the CPU oracle."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import software_renderer_b200 as srb  # noqa: E402
from software_renderer_b200 import scene as S  # noqa: E402
from software_renderer_b200.multi_gpu import gather_framebuffers, shard_range  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    NV, W, H = 13, 256, 192
    v, i = S.cube_sphere(40)
    views = S.fibonacci_views(NV)
    lo, hi = shard_range(NV, rank, world)
    r = srb.CudaRenderer(local)
    r.set_mesh(v, i)
    color = torch.empty((hi - lo, H, W, 4), dtype=torch.uint8, device=dev)
    depth = torch.empty((hi - lo, H, W), dtype=torch.float32, device=dev)
    if hi > lo:
        r.render_device(views[lo:hi], W, H, color.data_ptr(), depth.data_ptr(), stream=torch.cuda.current_stream().cuda_stream)
    full = gather_framebuffers(color, NV, dst=0)
    ok = True
    if rank == 0:
        from oracle import oracle as O
        got = full.cpu().numpy()
        for k in range(NV):
            rc, c, _, _ = O.render(v, i, views[k], W, H)
            ok &= rc == 0 and np.array_equal(got[k], c)
        print(f"nccl composite over {world} GPUs: {NV} views {W}x{H}, {'bit-exact vs oracle' if ok else 'MISMATCH'}", flush=True)
    r.close()
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
