#!/usr/bin/env python
"""Multi-GPU check, launched by tests/test_gather_gpu.py (or by hand) on a box with >= 2 GPUs:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 \
        tests/nccl_composite_check.py

Each rank renders its block of the view list on its own GPU in chunks (no collective on the data path); each finished
chunk is gathered to rank 0 through the C ABI (sr_gather: grouped ncclSend / ncclRecv on the context's gather stream)
while the next chunk renders; rank 0 compares every view bit for bit with the CPU oracle."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import software_renderer_b200 as srb  # noqa: E402
from software_renderer_b200 import scene as S  # noqa: E402
from software_renderer_b200.multi_gpu import gather_views, init_comm, shard_range  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("gloo")                     # host plumbing only (carries the NCCL id); the data path is sr_gather
    NCHUNK, PER, W, H = 3, 2 * world + 1, 256, 192      # uneven shards on purpose
    v, i = S.cube_sphere(40)
    views = S.fibonacci_views(NCHUNK * PER)
    r = srb.CudaRenderer(local)
    r.set_mesh(v, i)
    init_comm(r)
    stream = torch.cuda.Stream(device=dev)
    lo, hi = shard_range(PER, rank, world)
    n = hi - lo
    color = [torch.empty((max(n, 1), H, W, 4), dtype=torch.uint8, device=dev) for _ in range(2)]
    depth = [torch.empty((max(n, 1), H, W), dtype=torch.float32, device=dev) for _ in range(2)]
    full_c = torch.empty((NCHUNK, PER, H, W, 4), dtype=torch.uint8, device=dev) if rank == 0 else None
    full_d = torch.empty((NCHUNK, PER, H, W), dtype=torch.float32, device=dev) if rank == 0 else None
    for k in range(NCHUNK):
        b = k & 1
        r.gather_fence(2 * b, stream.cuda_stream)       # the gathers that read this buffer pair two chunks ago
        r.gather_fence(2 * b + 1, stream.cuda_stream)
        if n:
            r.render_device(views[k * PER + lo:k * PER + hi], W, H, color[b].data_ptr(), depth[b].data_ptr(),
                            stream=stream.cuda_stream, sync=False)
        gather_views(r, color[b][:n], PER, full_c[k] if rank == 0 else None, slot=2 * b, stream=stream.cuda_stream)
        gather_views(r, depth[b][:n], PER, full_d[k] if rank == 0 else None, slot=2 * b + 1, stream=stream.cuda_stream)
    r.sync()
    r.gather_sync()
    ok = True
    if rank == 0:
        from oracle import oracle as O
        got_c, got_d = full_c.cpu().numpy().reshape(-1, H, W, 4), full_d.cpu().numpy().reshape(-1, H, W)
        for k in range(NCHUNK * PER):
            rc, c, d, _ = O.render(v, i, views[k], W, H)
            ok &= rc == 0 and np.array_equal(got_c[k], c) and np.array_equal(got_d[k].view(np.uint32), d.view(np.uint32))
        print(f"sr_gather composite over {world} GPUs: {NCHUNK * PER} views {W}x{H} in {NCHUNK} overlapped chunks, "
              f"{'bit-exact vs oracle' if ok else 'MISMATCH'}", flush=True)
    r.comm_destroy()
    r.close()
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
