#!/usr/bin/env python
"""bench.py — frames/s of the rasterization hot path (BASELINE.json metric) on N B200s of one node.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload soup1m|ico80k|cs10m4k|views512] [--impl reference]
                  [--depth-test 1] [--split-frame 1]

A "step" is one render pass (Renderer.render(renderPass:), Renderer.swift:204-230) of the named workload with that
step's own transform.  Default workload: BASELINE.json's metric config — 1M-triangle soup at 1920x1080 (configs[2]).
  value  frames/s, whole job, mesh resident in HBM, images left in HBM; device time from CUDA events on the launching
         stream; L2 flushed between timed frames; max over ranks.
  e2e    same metric through the synchronous host-pointer C-ABI call (sr_render): per step the transform goes in and
         the colour and depth images come back to pinned host memory inside the timed region.
  --impl reference   the CPU restatement of the reference path (oracle/, the reference itself is Swift and cannot be
         built here) on all host cores, one bounded sample of the same workload per step.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (description, width, height)
    "soup1m": ("1M-triangle random soup, heavy overdraw, 1920x1080 (BASELINE.json configs[2])", 1920, 1080),
    "ico80k": ("icosphere subdiv 6 (81,920 tris), baked Lambert, 1920x1080 (configs[1])", 1920, 1080),
    "cs10m4k": ("cube-sphere n=913 (10,002,828 tris), 3840x2160 (configs[3])", 3840, 2160),
    "views512": ("cube-sphere n=289 (1,002,252 tris), 512x512 camera views (configs[4])", 512, 512),
}


def make_workload(name: str):
    from software_renderer_b200 import scene as S
    desc, W, H = WORKLOADS[name]
    if name == "soup1m":
        v, i = S.triangle_soup(1_000_000, W, H)
        xf = S.soup_transform
    elif name == "ico80k":
        v, i = S.icosphere(6)
        xf = lambda k: S.app_transform(k / 60.0)
    elif name == "cs10m4k":
        v, i = S.cube_sphere(913)
        xf = lambda k: S.app_transform(k / 60.0)
    else:
        v, i = S.cube_sphere(289)
        views = S.fibonacci_views(4096)
        xf = lambda k: views[k % 4096]
    return desc, W, H, v, i, xf


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons (NVML) during the timed region."""

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._halt = threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            pass

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {
            nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
            nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap",
        }
        while not self._halt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, nm in names.items():
                    if r & bit:
                        self.reasons.add(nm)
            except Exception:
                pass
            self._halt.wait(0.005)                 # the timed region of 100 frames is 15-50 ms

    def stop(self):
        self._halt.set()
        self.join(timeout=2)
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


def bind_to_gpu_numa_node(index: int) -> None:
    """N ranks on one box: run this rank on the CPUs next to its GPU, so that the pinned host images of the e2e leg
    (first touch) and the PCIe writes into them stay on the GPU's own socket.  Best effort."""
    try:
        import pynvml
        pynvml.nvmlInit()
        pynvml.nvmlDeviceSetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(index))
    except Exception:
        pass


def cpu_count() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def time_oracle(v, i, xf, W, H, nthreads: int, frames_per_thread: int, stride: int, depth_test: bool = False):
    """Runs the oracle on `nthreads` host threads, each rendering `frames_per_thread` samples (every `stride`-th
    triangle of the frame).  Returns seconds of wall time."""
    from oracle import oracle as O
    O.lib()
    idx = np.ascontiguousarray(i.reshape(-1, 3)[::stride].reshape(-1))
    bufs = [(np.empty((H, W, 4), np.uint8), np.empty((H, W), np.float32)) for _ in range(nthreads)]

    def work(t):
        for f in range(frames_per_thread):
            O.render(v, idx, xf(t * frames_per_thread + f), W, H, flags=O.DEPTH_LESS if depth_test else 0,
                     color_out=bufs[t][0], depth_out=bufs[t][1])

    th = [threading.Thread(target=work, args=(t,)) for t in range(nthreads)]
    t0 = time.perf_counter()
    for x in th:
        x.start()
    for x in th:
        x.join()
    return time.perf_counter() - t0


def run_reference(args):
    """Reference arm: the CPU restatement (oracle/) on all host cores; rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    desc, W, H, v, i, xf = make_workload(args.workload)
    ntris = i.size // 3
    cores = cpu_count()
    from oracle import oracle as O
    O.lib()
    # calibrate: one thread, a 1/64 sample of one frame
    stride0 = 64 if ntris >= 64_000 else 1
    t_cal = time_oracle(v, i, xf, W, H, 1, 1, stride0, bool(args.depth_test))
    full_frame_s = t_cal * stride0                       # estimate; cost is ~linear in triangles for these meshes
    budget = 150.0 / max(1, args.steps + args.warmup)    # seconds per step
    stride = 1
    while full_frame_s / stride > budget and ntris // (stride * 2) >= 1000:
        stride *= 2
    for _ in range(args.warmup):
        time_oracle(v, i, xf, W, H, cores, 1, stride, bool(args.depth_test))
    t = 0.0
    for _ in range(args.steps):
        t += time_oracle(v, i, xf, W, H, cores, 1, stride, bool(args.depth_test))
    # each step renders `cores` samples, each 1/stride of a frame's triangles (plus a full clear)
    frames = args.steps * cores / stride
    value = frames / t
    sample = (f"each step = {cores} host threads x 1 render pass of every {stride}-th triangle of the frame "
              f"({ntris // stride} of {ntris} tris, full {W}x{H} clear); frames counted as 1/{stride} each")
    line = {
        "impl": "reference", "metric": "frames_per_second", "value": value, "unit": "frames/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1000.0 * t / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32+i64", "data": "synthetic",
        "config": {"workload": desc, "width": W, "height": H, "triangles": ntris, "vertices": int(v.shape[0]),
                   "depth_test": bool(args.depth_test), "l2": "n/a (host cores)",
                   "parallelism": f"{cores} host threads, one render pass each per step"},
        "cpu_baseline": {"value": value, "unit": "frames/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "reference is Swift (no toolchain here): oracle/ C++ restatement of Renderer.swift, -O2 -ffp-contract=off",
    }
    print(json.dumps(line), flush=True)


def run_cuda(args):
    import torch
    import torch.distributed as dist
    import software_renderer_b200 as srb
    from software_renderer_b200 import scene as S
    from software_renderer_b200.multi_gpu import max_over_ranks

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        bind_to_gpu_numa_node(local)

    desc, W, H, v, i, xf = make_workload(args.workload)
    ntris, nverts = i.size // 3, int(v.shape[0])
    K, Wm = args.steps, args.warmup
    r = srb.CudaRenderer(local)
    r.depth_test = bool(args.depth_test)
    r.set_mesh(v, i)
    stream = torch.cuda.Stream(device=dev)
    color = torch.empty((H, W, 4), dtype=torch.uint8, device=dev)
    depth = torch.empty((H, W), dtype=torch.float32, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)      # > 126 MB L2
    # every rank renders its own frames: frame index = rank * (K + Wm) + step
    frame0 = rank * (K + Wm + 8)

    def frame(k):
        r.render_device(xf(frame0 + k), W, H, color.data_ptr(), depth.data_ptr(), stream=stream.cuda_stream, sync=False)

    with torch.cuda.stream(stream):
        frame(0)
        r.sync()                                        # lets the adaptive batching policy settle
        for k in range(1, Wm):
            frame(k)
        r.sync()
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        sampler = ClockSampler(local)
        sampler.start()
        t_wall0 = time.perf_counter()
        launches = 0
        for k in range(K):
            if not os.environ.get("SR_BENCH_NOFLUSH"):
                flush.fill_(k & 0xFF)                   # L2 flush between timed frames (not timed)
            evs[k][0].record(stream)
            frame(Wm + k)                               # asynchronous; calls accumulate, no host round trip
            evs[k][1].record(stream)
        r.sync()                                        # one status read-back for all K frames
        launches = r.stats()["gpu_launches"]
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        t_wall = time.perf_counter() - t_wall0
        clocks = sampler.stop()
    per_frame = np.array([a.elapsed_time(b) for a, b in evs])         # this rank's frames, ms each
    dev_ms = float(per_frame.sum())
    dev_ms = max_over_ranks(dev_ms, device=dev)
    value = world * K / (dev_ms / 1000.0)

    # ---- e2e: host-pointer C-ABI calls, pinned host images; H2D of the transform and D2H of both images are inside
    # the timed region.  (a) sr_render: one synchronous call per frame, the strict drop-in for render(renderPass:);
    # (b) sr_render_batch: the same frames handed over 16 at a time, D2H of frame i overlapped with frame i+1.
    hc = torch.empty((H, W, 4), dtype=torch.uint8).pin_memory()
    hd = torch.empty((H, W), dtype=torch.float32).pin_memory()
    rp = srb.RenderPass(srb.Image(hc.numpy(), W, H), srb.Image(hd.numpy(), W, H), v, i, srb.PrimitiveType.triangle, None)
    r._mesh_key = (v, i)                                # mesh already uploaded (static across frames)
    for k in range(max(3, Wm)):
        rp.transform = xf(frame0 + k)
        r.render(rp)

    def timed_passes(one_pass):
        """The K e2e frames are repeated until the wall-clock region is ~0.3 s (a 1080p frame takes 0.3 ms, a 512x512
        view 0.05 ms: K = 100 frames alone would be a region of a few ms, at the mercy of one scheduler hiccup on
        one of N ranks).  Returns whole-job frames/s (max over ranks of the wall time) and the pass count."""
        t0 = time.perf_counter()
        one_pass()
        passes = int(min(200, max(1, np.ceil(0.3 / max(time.perf_counter() - t0, 1e-4)))))
        passes = int(max_over_ranks(float(passes), device=dev))          # the same count on every rank
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        for _ in range(passes):
            one_pass()
        t = max_over_ranks(time.perf_counter() - t0, device=dev)
        return world * K * passes / t, passes

    def sync_pass():
        for k in range(K):
            rp.transform = xf(frame0 + Wm + k)
            r.render(rp)

    e2e_sync, sync_passes = timed_passes(sync_pass)
    CH = 16
    bc = torch.empty((CH, H, W, 4), dtype=torch.uint8).pin_memory()
    bd = torch.empty((CH, H, W), dtype=torch.float32).pin_memory()
    mats = np.stack([xf(frame0 + Wm + k) for k in range(K)])
    r.render_views(mats[:min(K, CH)], W, H, color_out=bc.numpy(), depth_out=bd.numpy())      # warm-up

    def batch_pass():
        for s0 in range(0, K, CH):
            n = min(CH, K - s0)
            r.render_views(mats[s0:s0 + n], W, H, color_out=bc.numpy()[:n], depth_out=bd.numpy()[:n])

    e2e_value, batch_passes = timed_passes(batch_pass)

    line = None
    if rank == 0:
        # per-kernel split of one frame (extra events between kernels; separate from the timed loop)
        r.set_profiling(True)
        acc = {}
        reps = 5
        for k in range(reps):
            flush.fill_(1)
            torch.cuda.synchronize(dev)
            r.render_device(xf(frame0 + k), W, H, color.data_ptr(), depth.data_ptr(), stream=stream.cuda_stream)
            for nm, ms in r.kernel_times().items():
                acc[nm] = acc.get(nm, 0.0) + ms / reps
        r.set_profiling(False)
        stats = r.stats()
        hbm, peak_kind = peaks()
        # algorithmic bytes: indices 12 B/tri, vertex xyz+rgb 24 B, colour+depth 8 B/pixel, split over
        # the kernels that touch them
        alg = {"k_transform": 12 * nverts, "k_setup": 12 * ntris, "k_tile": 12 * nverts + 8 * W * H,
               "k_resolve": 12 * nverts + 8 * W * H}
        top = max(acc, key=acc.get) if acc else "k_tile"
        top_ms = acc.get(top, float("nan"))
        top_bytes = alg.get(top, 0)
        achieved = top_bytes / (top_ms * 1e-3) / 1e9 if top_ms and top_ms == top_ms else float("nan")
        traffic = None
        tp = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tp):
            try:
                traffic = json.load(open(tp)).get(args.workload + ("_depth" if args.depth_test else ""), {}).get(top)
            except Exception:
                traffic = None
        frame_bytes = S.algorithmic_bytes(ntris, nverts, W, H)
        frame_ms = dev_ms / K
        # CPU baseline: the oracle on one host core (the reference loop is sequential, Renderer.swift:222), about 10 s
        # of CPU work: calibrate on a 1/8 sample, then whole frames (or the largest 1/stride sample that fits)
        stride0 = 8 if ntris >= 500_000 else 1
        t_cal = time_oracle(v, i, xf, W, H, 1, 1, stride0, bool(args.depth_test))
        full_s = t_cal * stride0
        stride = 1
        while full_s / stride > 15.0:
            stride *= 2
        nframes = int(max(1, min(2000, 10.0 / max(full_s / stride, 1e-4))))
        t_cpu = time_oracle(v, i, xf, W, H, 1, nframes, stride, bool(args.depth_test))
        cpu_value = (nframes / stride) / t_cpu
        line = {
            "metric": "frames_per_second", "value": value, "unit": "frames/s", "n_gpus": world, "steps": K, "warmup": Wm,
            "ms_per_step": frame_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32+i64", "data": "synthetic",
            "config": {"workload": desc, "width": W, "height": H, "triangles": ntris, "vertices": nverts,
                       "depth_test": bool(args.depth_test), "l2": "flushed between timed frames (256 MiB fill)",
                       "parallelism": f"{world} independent view streams, one per GPU"},
            "gtri_per_s": value * ntris / 1e9,
            "frame_ms_rank0": {"median": float(np.median(per_frame)), "p10": float(np.percentile(per_frame, 10)),
                               "p90": float(np.percentile(per_frame, 90))},
            "frame_roofline": {"bound": "hbm", "achieved": frame_bytes / (frame_ms * 1e-3) / 1e9, "peak": hbm,
                               "unit": "GB/s", "frac": frame_bytes / (frame_ms * 1e-3) / 1e9 / hbm,
                               "bytes_per_frame": frame_bytes, "peak_kind": peak_kind},
            "roofline": {"bound": "hbm", "kernel": top, "achieved": achieved, "peak": hbm, "unit": "GB/s",
                         "frac": achieved / hbm, "traffic": traffic, "algorithmic_bytes": top_bytes,
                         "kernel_ms": top_ms, "peak_kind": peak_kind},
            "kernel_ms": acc,
            "cpu_baseline": {"value": cpu_value, "unit": "frames/s", "cores": 1, "kind": "port",
                             "sample": f"{nframes} render pass(es) of every {stride}-th triangle ({ntris // stride} of {ntris} "
                                       f"tris, full clear) on 1 core, {t_cpu:.2f} s in total, each pass counted as 1/{stride} "
                                       f"frame; host has {cpu_count()} cores"},
            "e2e": {"value": e2e_value, "unit": "frames/s", "h2d_bytes_per_step": 64, "d2h_bytes_per_step": 8 * W * H,
                    "api": "sr_render_batch, 16 frames per call, pinned host images, mesh uploaded once by sr_set_mesh",
                    "frames_timed_per_rank": K * batch_passes,
                    "sync_per_frame": {"value": e2e_sync, "unit": "frames/s", "api": "sr_render, one synchronous call per frame",
                                       "frames_timed_per_rank": K * sync_passes}},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "wall_s_timed_region": t_wall,
            "stats_last_frame": {k: (int(x) if k != "gpu_ms" else float(x)) for k, x in stats.items()},
        }
    r.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if line is not None:
        print(json.dumps(line), flush=True)


def run_split(args):
    """--split-frame 1: ONE frame per step, split by rows over the ranks (sr_set_band).  Each
    rank renders its band of tile rows of the same frame, the bands are gathered to rank 0 over NCCL inside the timed
    region (this mode's only collective), so value = whole frames/s as seen by rank 0: strong scaling.  Not the
    headline mode (independent frames need no exchange at all); it measures single-frame latency across GPUs."""
    import torch
    import torch.distributed as dist
    import software_renderer_b200 as srb
    from software_renderer_b200.multi_gpu import band_range, gather_bands, max_over_ranks

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    desc, W, H, v, i, xf = make_workload(args.workload)
    ntris, nverts = i.size // 3, int(v.shape[0])
    K, Wm = args.steps, args.warmup
    r = srb.CudaRenderer(local)
    r.depth_test = bool(args.depth_test)
    r.set_mesh(v, i)
    y0, y1 = band_range(H, rank, world)
    assert y1 > y0, "fewer tile rows than ranks"
    if world > 1:
        r.set_band(y0, y1)
    color = torch.empty((y1 - y0, W, 4), dtype=torch.uint8, device=dev)
    depth = torch.empty((y1 - y0, W), dtype=torch.float32, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    stream = torch.cuda.Stream(device=dev)
    full = None

    def frame(k):
        r.render_device(xf(k), W, H, color.data_ptr(), depth.data_ptr(), stream=stream.cuda_stream, sync=False)
        if world > 1:
            return gather_bands(color, H, dst=0), gather_bands(depth, H, dst=0)
        return color, depth

    with torch.cuda.stream(stream):
        frame(0)
        r.sync()
        for k in range(1, Wm):
            frame(k)
        r.sync()
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        sampler = ClockSampler(local)
        sampler.start()
        for k in range(K):
            flush.fill_(k & 0xFF)
            evs[k][0].record(stream)
            full = frame(Wm + k)
            evs[k][1].record(stream)
        r.sync()
        launches = r.stats()["gpu_launches"]
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        clocks = sampler.stop()
    per_frame = np.array([a.elapsed_time(b) for a, b in evs])
    dev_ms = max_over_ranks(float(per_frame.sum()), device=dev)
    value = K / (dev_ms / 1000.0)
    line = None
    if rank == 0:
        # the stitched frame of the last step against this rank's own whole-frame render of the same transform
        got_c, got_d = full[0].cpu().numpy(), full[1].cpu().numpy()
        r.set_band()
        want_c, want_d = r.render_views(xf(Wm + K - 1)[None], W, H)
        same = bool(np.array_equal(got_c, want_c[0]) and np.array_equal(got_d.view(np.uint32), want_d[0].view(np.uint32)))
        hbm, peak_kind = peaks()
        from software_renderer_b200 import scene as S
        frame_bytes = S.algorithmic_bytes(ntris, nverts, W, H)
        line = {
            "metric": "frames_per_second", "value": value, "unit": "frames/s", "n_gpus": world, "steps": K, "warmup": Wm,
            "ms_per_step": dev_ms / K, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32+i64", "data": "synthetic",
            "config": {"workload": desc, "width": W, "height": H, "triangles": ntris, "vertices": nverts,
                       "depth_test": bool(args.depth_test), "l2": "flushed between timed frames (256 MiB fill)",
                       "parallelism": f"one frame split into {world} bands of tile rows, NCCL gather of the bands to rank 0 inside the timed region"},
            "frame_ms_rank0": {"median": float(np.median(per_frame)), "p10": float(np.percentile(per_frame, 10)),
                               "p90": float(np.percentile(per_frame, 90))},
            "frame_roofline": {"bound": "hbm", "achieved": frame_bytes / (dev_ms / K * 1e-3) / 1e9, "peak": hbm, "unit": "GB/s",
                               "frac": frame_bytes / (dev_ms / K * 1e-3) / 1e9 / hbm, "bytes_per_frame": frame_bytes, "peak_kind": peak_kind},
            "bands_equal_whole_frame": same,
            "gathered_bytes_per_frame": 0 if world == 1 else 8 * W * (H - (y1 - y0)),
            "gpu_launches": int(launches), "clocks": clocks,
        }
    r.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if line is not None:
        print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="cuda", choices=["cuda", "reference"])
    ap.add_argument("--workload", default="soup1m", choices=sorted(WORKLOADS))
    ap.add_argument("--depth-test", type=int, default=0)
    ap.add_argument("--split-frame", type=int, default=0, help="1: every step is ONE frame split by rows over the ranks")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "cuda" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    elif args.split_frame:
        run_split(args)
    else:
        run_cuda(args)


if __name__ == "__main__":
    main()
