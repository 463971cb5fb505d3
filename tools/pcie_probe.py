#!/usr/bin/env python
"""Bare device->host copy probe: the roofline of bench.py's e2e leg ("Next round" item 1a).

N processes, one per GPU, each issues back-to-back cudaMemcpyAsync D2H copies of one frame-sized block (default
16,588,800 B = colour + depth of a 1080p frame; also the colour plane alone) into (a) pinned and (b) pageable host
memory, all ranks started together behind a barrier.  Prints one JSON line per N with GB/s per GPU and aggregate,
and what that means in 1080p frames/s.  No kernel of this repo runs: this is what the box can deliver.

  python tools/pcie_probe.py --sweep 1,2,4,8 [--bytes 16588800] [--reps 200] > gpurun_out/pcie_probe.jsonl
"""
import argparse
import json
import sys
import time

import torch
import torch.multiprocessing as mp


def worker(rank, world, nbytes, reps, barrier, out_q):
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    try:
        import pynvml
        pynvml.nvmlInit()
        pynvml.nvmlDeviceSetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(rank))
    except Exception:
        pass
    src = torch.empty(nbytes, dtype=torch.uint8, device=dev).fill_(rank + 1)
    pinned = [torch.empty(nbytes, dtype=torch.uint8).pin_memory() for _ in range(2)]
    pageable = [torch.empty(nbytes, dtype=torch.uint8) for _ in range(2)]
    for b in pageable:
        b.fill_(0)                                   # touch the pages
    st = torch.cuda.Stream(device=dev)
    res = {}
    for name, dst in (("pinned", pinned), ("pageable", pageable)):
        n = reps if name == "pinned" else max(8, reps // 8)
        with torch.cuda.stream(st):
            for k in range(3):
                dst[k & 1].copy_(src, non_blocking=True)
            st.synchronize()
            barrier.wait()
            t0 = time.perf_counter()
            for k in range(n):
                dst[k & 1].copy_(src, non_blocking=True)
            st.synchronize()
            t = time.perf_counter() - t0
        barrier.wait()
        res[name] = nbytes * n / t / 1e9
    # host -> device of the same block (the strict drop-in call re-sends the mesh every frame)
    with torch.cuda.stream(st):
        for k in range(3):
            src.copy_(pinned[0], non_blocking=True)
        st.synchronize()
        barrier.wait()
        t0 = time.perf_counter()
        for k in range(reps):
            src.copy_(pinned[k & 1], non_blocking=True)
        st.synchronize()
        res["h2d_pinned"] = nbytes * reps / (time.perf_counter() - t0) / 1e9
    barrier.wait()
    out_q.put((rank, res))


def run(world, nbytes, reps):
    ctx = mp.get_context("spawn")
    barrier = ctx.Barrier(world)
    q = ctx.Queue()
    ps = [ctx.Process(target=worker, args=(r, world, nbytes, reps, barrier, q)) for r in range(world)]
    for p in ps:
        p.start()
    got = dict(q.get() for _ in range(world))
    for p in ps:
        p.join()
    line = {"probe": "bare cudaMemcpyAsync", "n_gpus": world, "block_bytes": nbytes, "reps": reps}
    for key in ("pinned", "pageable", "h2d_pinned"):
        per = [got[r][key] for r in range(world)]
        line[key] = {"gbs_per_gpu": [round(x, 2) for x in per], "gbs_aggregate": round(sum(per), 2),
                     "blocks_per_s_aggregate": round(sum(per) * 1e9 / nbytes, 1)}
    return line


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sweep", default="1")
    ap.add_argument("--bytes", type=int, default=8 * 1920 * 1080)
    ap.add_argument("--reps", type=int, default=200)
    a = ap.parse_args()
    have = torch.cuda.device_count()
    for n in [int(x) for x in a.sweep.split(",")]:
        if n > have:
            print(json.dumps({"probe": "skipped", "n_gpus": n, "why": f"only {have} GPU(s) visible"}), flush=True)
            continue
        for nbytes in (a.bytes, a.bytes // 2):
            print(json.dumps(run(n, nbytes, a.reps)), flush=True)


if __name__ == "__main__":
    sys.exit(main())
