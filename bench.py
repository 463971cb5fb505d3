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


class OneLineStdout:
    """stdout carries exactly ONE JSON line.  Libraries print there from C code (NCCL's "NCCL version ..." banner ignores
    NCCL_DEBUG_FILE): for the duration of the run file descriptor 1 points at stderr, and emit() writes the line to the
    descriptor that was stdout."""

    def __init__(self):
        sys.stdout.flush()
        self.real = os.dup(1)
        os.dup2(2, 1)

    def emit(self, line: str):
        sys.stdout.flush()
        os.write(self.real, (line + "\n").encode())


OUT = None

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
        "config": workload_config(desc, W, H, ntris, int(v.shape[0]), args.depth_test),
        "run": {"l2": "n/a (host cores)", "parallelism": f"{cores} host threads, one render pass each per step"},
        "cpu_baseline": {"value": value, "unit": "frames/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "reference is Swift (no toolchain here): oracle/ C++ restatement of Renderer.swift, -O2 -ffp-contract=off",
    }
    OUT.emit(json.dumps(line))


def workload_config(desc, W, H, ntris, nverts, depth_test):
    """`config` of the JSON line: the workload and nothing else, identical in both arms (run-mode facts go to `run`)."""
    return {"workload": desc, "width": W, "height": H, "triangles": ntris, "vertices": nverts, "depth_test": bool(depth_test)}


REGION_STATS = {}


def device_fps(r, xf, W, H, color, depth, stream, flush, K, Wm, frame0, dev):
    """K timed frames, device resident, L2 flushed between frames; returns (per-frame ms array, launches, wall s)."""
    import torch

    # matrices converted up front: the enqueue loop must stay ahead of a 0.06 - 0.2 ms frame (with numpy building a
    # matrix per iteration it did not: 0.43 ms of host time per frame, the kernels of a frame were submitted just in time
    # and programmatic dependent launch had nothing to overlap)
    mats = [r.colmajor(xf(frame0 + k)) for k in range(Wm + K)]
    cptr, dptr, sptr = color.data_ptr(), depth.data_ptr(), stream.cuda_stream

    def frame(k):
        r.render_device(mats[k], W, H, cptr, dptr, stream=sptr, sync=False, colmajor=True)

    with torch.cuda.stream(stream):
        frame(0)
        r.sync()                                        # lets the adaptive launch policies settle
        for k in range(1, Wm):
            frame(k)
        r.sync()
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
        torch.cuda.synchronize(dev)
        t0 = time.perf_counter()
        for k in range(K):
            if not os.environ.get("SR_BENCH_NOFLUSH"):
                flush.fill_(k & 0xFF)                   # L2 flush between timed frames (not timed)
            evs[k][0].record(stream)
            frame(Wm + k)                               # asynchronous; calls accumulate, no host round trip
            evs[k][1].record(stream)
        r.sync()                                        # one status read-back for all K frames
        ev_tail = torch.cuda.Event(enable_timing=True)
        ev_tail.record(stream)
        region = r.stats()                              # counters of the K timed frames (one accumulated call)
        launches = region["gpu_launches"]
        torch.cuda.synchronize(dev)
        wall = time.perf_counter() - t0
    # frames re-run after a failed speculation (list capacity, no-binning guess) run at the sync, outside
    # the per-frame event pairs: they are reported so that a line that lived on re-runs cannot hide it
    REGION_STATS.clear()
    REGION_STATS.update({k: int(region[k]) for k in ("bin_regrows", "nobin_reruns", "tile_pairs", "pairs_consumed", "depth_culled")})
    per_frame = np.array([a.elapsed_time(b) for a, b in evs])
    reruns = REGION_STATS["bin_regrows"] + REGION_STATS["nobin_reruns"]
    if reruns:                                          # charge the re-runs (and the host round trip before them) to the last frame
        torch.cuda.synchronize(dev)
        per_frame[-1] += evs[-1][1].elapsed_time(ev_tail)
        REGION_STATS["rerun_tail_ms_charged"] = float(evs[-1][1].elapsed_time(ev_tail))
    return per_frame, int(launches), wall


def kernel_split(r, xf, W, H, color, depth, stream, flush, frame0, dev, reps=5):
    """Per-kernel device time of one frame (extra events between kernels; separate from the timed loop)."""
    import torch
    r.set_profiling(True)
    acc = {}
    with torch.cuda.stream(stream):
        for k in range(reps):
            flush.fill_(1)
            torch.cuda.synchronize(dev)
            r.render_device(xf(frame0 + k), W, H, color.data_ptr(), depth.data_ptr(), stream=stream.cuda_stream)
            for nm, ms in r.kernel_times().items():
                acc[nm] = acc.get(nm, 0.0) + ms / reps
    r.set_profiling(False)
    return acc


def roofline_of(acc, ntris, nverts, W, H, workload_key):
    """Roofline line of the dominant kernel.  Algorithmic bytes (12 B per triangle of indices, 24 B per
    vertex of xyz + rgb, 8 B per pixel of colour + depth) are split over the kernels that touch them:
    positions are read by k_transform, or - in batched frames, which do not launch it - by k_setup."""
    hbm, peak_kind = peaks()
    # (with the box pre-pass of batched frames the later batches' indices and positions are read by k_late_boxes; charged in
    # full to either kernel, an upper bound for each: the split depends on the batch sizes)
    alg = {"k_transform": 12 * nverts, "k_setup": 12 * ntris + (0 if "k_transform" in acc else 12 * nverts),
           "k_late_boxes": 12 * ntris + 12 * nverts,
           "k_tile": 12 * nverts + 8 * W * H, "k_resolve": 12 * nverts + 8 * W * H}
    top = max(acc, key=acc.get) if acc else "k_tile"
    top_ms = acc.get(top, float("nan"))
    top_bytes = alg.get(top, 0)
    achieved = top_bytes / (top_ms * 1e-3) / 1e9 if top_ms and top_ms == top_ms else float("nan")
    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        try:
            traffic = json.load(open(tp)).get(workload_key, {}).get(top)
        except Exception:
            traffic = None
    return {"bound": "hbm", "kernel": top, "achieved": achieved, "peak": hbm, "unit": "GB/s", "frac": achieved / hbm,
            "traffic": traffic, "algorithmic_bytes": top_bytes, "kernel_ms": top_ms, "peak_kind": peak_kind}


def frame_roofline_of(frame_ms, ntris, nverts, W, H):
    from software_renderer_b200 import scene as S
    hbm, peak_kind = peaks()
    b = S.algorithmic_bytes(ntris, nverts, W, H)
    return {"bound": "hbm", "achieved": b / (frame_ms * 1e-3) / 1e9, "peak": hbm, "unit": "GB/s",
            "frac": b / (frame_ms * 1e-3) / 1e9 / hbm, "bytes_per_frame": b, "peak_kind": peak_kind}


class E2E:
    """End-to-end legs: the host-pointer C-ABI calls, host<->device copies inside the timed wall-clock region."""

    def __init__(self, world, dev, K):
        self.world, self.dev, self.K = world, dev, K

    def timed(self, one_pass, frames_per_pass):
        """`one_pass` is repeated until the wall-clock region is ~0.3 s (a 1080p frame takes 0.2 ms: K frames alone would be
        a region of a few ms, at the mercy of one scheduler hiccup on one of N ranks).  Whole-job frames/s, max over ranks."""
        import torch.distributed as dist
        from software_renderer_b200.multi_gpu import max_over_ranks
        t0 = time.perf_counter()
        one_pass()
        passes = int(min(200, max(1, np.ceil(0.3 / max(time.perf_counter() - t0, 1e-4)))))
        passes = int(max_over_ranks(float(passes), device=self.dev))          # the same count on every rank
        if self.world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        for _ in range(passes):
            one_pass()
        t = max_over_ranks(time.perf_counter() - t0, device=self.dev)
        return self.world * frames_per_pass * passes / t, frames_per_pass * passes


def e2e_suite(r, local, world, dev, v, i, xf, W, H, K, Wm, frame0, full: bool):
    """Returns the `e2e` object.  Headline (`value`): sr_render_batch, 16 frames per call, page-locked caller images, mesh
    uploaded once.  `full` (rank 0's N = 1 run) adds the variants a drop-in caller can meet: pageable images, the strict
    per-call mesh hand-over of the reference (App.swift:163), and the three ways to treat the constant depth plane."""
    import torch
    import software_renderer_b200 as srb
    e = E2E(world, dev, K)
    CH = 16
    mats = np.stack([xf(frame0 + Wm + k) for k in range(K)])

    def batch_leg(rr, bc, bd):
        rr.render_views(mats[:min(K, CH)], W, H, color_out=bc, depth_out=bd, want_depth=bd is not None)      # warm-up

        def one():
            for s0 in range(0, K, CH):
                n = min(CH, K - s0)
                rr.render_views(mats[s0:s0 + n], W, H, color_out=bc[:n], depth_out=None if bd is None else bd[:n], want_depth=bd is not None)
        return e.timed(one, K)

    bc = r.host_array((CH, H, W, 4), np.uint8)          # sr_host_alloc: page-locked
    bd = r.host_array((CH, H, W), np.float32)
    value, frames = batch_leg(r, bc, bd)
    depth_mode = r.depth_test
    out = {"value": value, "unit": "frames/s", "h2d_bytes_per_step": 64,
           "d2h_bytes_per_step": (8 if depth_mode else 4) * W * H,
           "api": "sr_render_batch, 16 frames per call, page-locked host images (sr_host_alloc), mesh uploaded once by sr_set_mesh; "
                  + ("colour and depth cross PCIe" if depth_mode else
                     "colour crosses PCIe, the constant +inf depth plane (depth test off, Renderer.swift:206) is written into the caller's image by the library's host threads"),
           "frames_timed_per_rank": frames}
    # the same call with the depth plane left alone on the caller's word / not asked for at all
    if not depth_mode:
        r.depth_untouched = True
        out["depth_untouched"] = {"value": batch_leg(r, bc, bd)[0], "unit": "frames/s", "api": "same, SR_FLAG_DEPTH_UNTOUCHED: the caller's depth image still holds the plane of the previous call"}
        r.depth_untouched = False
    out["no_depth_out"] = {"value": batch_leg(r, bc, None)[0], "unit": "frames/s", "api": "same, SR_FLAG_NO_DEPTH_OUT (colour only)"}
    # bare copies of the same bytes on every rank at once: what this box can deliver (the e2e roofline)
    out["pcie_probe"] = pcie_probe(e, local, dev, W, H, bc, bd, r)
    if not full:
        return out
    if not depth_mode:
        os.environ["SR_FILL_THREADS"] = "0"
        with srb.CudaRenderer(local) as r0:
            r0.set_mesh(v, i)
            out["depth_over_pcie"] = {"value": batch_leg(r0, bc, bd)[0], "unit": "frames/s", "d2h_bytes_per_step": 8 * W * H,
                                      "api": "same, SR_FILL_THREADS=0: the +inf plane is produced on the device and copied (round-1 behaviour)"}
        del os.environ["SR_FILL_THREADS"]
    pc, pd = np.empty((CH, H, W, 4), np.uint8), np.empty((CH, H, W), np.float32)
    out["pageable_images"] = {"value": batch_leg(r, pc, pd)[0], "unit": "frames/s", "api": "same call, ordinary (pageable) caller images, as App.swift:59-60 allocates them"}
    # one synchronous call per frame, static mesh (RenderPass.mesh_version)
    rp = srb.RenderPass(srb.Image(bc[0], W, H), srb.Image(bd[0], W, H), v, i, srb.PrimitiveType.triangle, None, mesh_version=1)

    def sync_pass():
        for k in range(K):
            rp.transform = mats[k]
            r.render(rp)
    sync_pass()
    out["sync_per_frame"] = {"value": e.timed(sync_pass, K)[0], "unit": "frames/s", "api": "sr_render, one synchronous call per frame, page-locked images, static mesh (mesh_version)"}
    # the reference's own contract: vertices and indices travel with EVERY call (App.swift:163) -> sr_render_pass
    KS = max(4, min(K, 16))
    mesh_bytes = int(v.nbytes + np.asarray(i, np.int64).nbytes)
    pv = r.host_array(v.shape, np.float32)
    pi = r.host_array((i.size,), np.int64)
    pv[:] = v
    pi[:] = i
    strict = {}
    for name, (vv, ii, cc, dd) in {"page_locked": (pv, pi, bc[0], bd[0]), "pageable": (v, np.ascontiguousarray(i, np.int64), pc[0], pd[0])}.items():
        rps = srb.RenderPass(srb.Image(cc, W, H), srb.Image(dd, W, H), vv, ii, srb.PrimitiveType.triangle, None)

        def strict_pass():
            for k in range(KS):
                rps.transform = mats[k % K]
                r.render(rps)
        strict_pass()
        strict[name] = e.timed(strict_pass, KS)[0]
    out["strict_render_pass"] = {"value": strict["pageable"], "unit": "frames/s", "page_locked_arrays": strict["page_locked"],
                                 "h2d_bytes_per_step": mesh_bytes + 64,
                                 "api": "sr_render_pass: mesh arrays (Swift layout: 32-B vertices, Int indices) re-sent on every call, as the reference hands them over; "
                                        "value = ordinary caller memory for arrays and images, page_locked_arrays = both from sr_host_alloc"}
    r.set_mesh(v, i)
    return out


def pcie_probe(e, local, dev, W, H, bc, bd, r):
    """Bare cudaMemcpyAsync D2H of one frame's colour plane (and colour + depth) into the same page-locked images, all ranks at
    once: GB/s per GPU and the frames/s that bandwidth allows.  No kernel of this repo runs.  `colour_plus_host_fill` is the
    ceiling of the DEFAULT e2e contract with the depth test off: the colour plane crosses PCIe while the library's host
    threads write the +inf depth plane (sr_host_fill_depth) - 8 bytes per pixel of fresh data land in host memory per frame."""
    import torch
    src = torch.empty(8 * W * H, dtype=torch.uint8, device=dev)
    dst_c = torch.from_numpy(bc.reshape(-1))
    dst_d = torch.from_numpy(bd.reshape(-1).view(np.uint8))
    st = torch.cuda.Stream(device=dev)
    res = {}
    for name, planes in (("colour", 1), ("colour_and_depth", 2)):
        n = 64

        def one():
            with torch.cuda.stream(st):
                for k in range(n):
                    o = (k % 16) * 4 * W * H
                    dst_c[o:o + 4 * W * H].copy_(src[:4 * W * H], non_blocking=True)
                    if planes == 2:
                        dst_d[o:o + 4 * W * H].copy_(src[4 * W * H:], non_blocking=True)
                st.synchronize()
        fps, _ = e.timed(one, n)
        res[name] = {"frames_per_s": fps, "gbs_aggregate": fps * planes * 4 * W * H / 1e9}

    def fill_one():
        n = 64
        with torch.cuda.stream(st):
            for k in range(n):
                o = (k % 16) * 4 * W * H
                dst_c[o:o + 4 * W * H].copy_(src[:4 * W * H], non_blocking=True)
                r.host_fill_depth(bd[k % 16])
            st.synchronize()
    fps, _ = e.timed(fill_one, 64)
    res["colour_plus_host_fill"] = {"frames_per_s": fps, "host_bytes_written_gbs": fps * 8 * W * H / 1e9}
    return res


def parity_check(r, v, i, m, W, H, depth_test):
    """One frame of the product path against the oracle (bit-exact colour and depth).  The oracle is the checker here."""
    from oracle import oracle as O
    c, d = r.render_views(np.asarray(m)[None], W, H)
    rc, oc, od, _ = O.render(v, i, m, W, H, flags=O.DEPTH_LESS if depth_test else 0)
    return bool(rc == 0 and np.array_equal(c[0], oc) and np.array_equal(d[0].view(np.uint32), od.view(np.uint32)))


def workload_record(name, depth_test, local, dev, stream, flush, K=20, Wm=5):
    """One entry of the `workloads` sub-record: device-resident frames/s, frame roofline, batch e2e and a parity bit."""
    import torch
    import software_renderer_b200 as srb
    desc, W, H, v, i, xf = make_workload(name)
    ntris, nverts = i.size // 3, int(v.shape[0])
    with srb.CudaRenderer(local) as r:
        r.depth_test = bool(depth_test)
        r.set_mesh(v, i)
        color = torch.empty((H, W, 4), dtype=torch.uint8, device=dev)
        depth = torch.empty((H, W), dtype=torch.float32, device=dev)
        per_frame, launches, _ = device_fps(r, xf, W, H, color, depth, stream, flush, K, Wm, 0, dev)
        region_stats = dict(REGION_STATS)
        acc = kernel_split(r, xf, W, H, color, depth, stream, flush, 0, dev, reps=3)
        frame_ms = float(per_frame.mean())
        CH = 8
        bc = r.host_array((CH, H, W, 4), np.uint8)
        bd = r.host_array((CH, H, W), np.float32)
        mats = np.stack([xf(Wm + k) for k in range(CH)])
        r.render_views(mats, W, H, color_out=bc, depth_out=bd)
        t0 = time.perf_counter()
        reps = 0
        while time.perf_counter() - t0 < 0.25:
            r.render_views(mats, W, H, color_out=bc, depth_out=bd)
            reps += 1
        e2e = CH * reps / (time.perf_counter() - t0)
        ok = parity_check(r, v, i, xf(3), W, H, depth_test)
        stats = r.stats()
    return {"config": workload_config(desc, W, H, ntris, nverts, depth_test), "value": 1000.0 / frame_ms, "unit": "frames/s",
            "ms_per_step": frame_ms, "steps": K, "gtri_per_s": ntris / frame_ms / 1e6,
            "frame_roofline": frame_roofline_of(frame_ms, ntris, nverts, W, H),
            "roofline": roofline_of(acc, ntris, nverts, W, H, name + ("_depth" if depth_test else "")),
            "kernel_ms": acc, "e2e": {"value": e2e, "unit": "frames/s", "api": "sr_render_batch, 8 frames per call, page-locked images"},
            "parity_vs_oracle": ok, "gpu_launches_per_frame": launches // K, "timed_region_stats": region_stats,
            "stats_last_frame": {k: (int(x) if k != "gpu_ms" else float(x)) for k, x in stats.items()}}


def cfg5_leg(local, rank, world, dev, nviews=4096, per_rank_chunk=128):
    """BASELINE.json configs[4]: 4096 camera views of the 1M-triangle cube-sphere at 512x512, sharded over the ranks (no
    exchange while rendering), (i) left on the GPUs, (ii) gathered to rank 0 through sr_gather (NCCL send/recv on the
    context's gather stream) with the gather of chunk k overlapping the rendering of chunk k+1.  Device time, max over ranks."""
    import torch
    import torch.distributed as dist
    import software_renderer_b200 as srb
    from software_renderer_b200.multi_gpu import gather_views, init_comm, max_over_ranks, shard_range
    desc, W, H, v, i, xf = make_workload("views512")
    from software_renderer_b200 import scene as S
    views = S.fibonacci_views(nviews)
    ntris, nverts = i.size // 3, int(v.shape[0])
    C = per_rank_chunk * world                           # views per chunk, all ranks together
    nchunks = (nviews + C - 1) // C
    stream = torch.cuda.Stream(device=dev)
    with srb.CudaRenderer(local) as r:
        r.set_mesh(v, i)
        if world > 1:
            init_comm(r)
        color = [torch.empty((per_rank_chunk, H, W, 4), dtype=torch.uint8, device=dev) for _ in range(2)]
        depth = [torch.empty((per_rank_chunk, H, W), dtype=torch.float32, device=dev) for _ in range(2)]
        full_c = torch.empty((nviews, H, W, 4), dtype=torch.uint8, device=dev) if rank == 0 else None
        full_d = torch.empty((nviews, H, W), dtype=torch.float32, device=dev) if rank == 0 else None

        inf_plane = float("inf")

        def run(gather: bool, render: bool = True):
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize(dev)
            if world > 1:
                dist.barrier()
            with torch.cuda.stream(stream):
                ev0.record(stream)
                for k in range(nchunks):
                    base, n_all = k * C, min(C, nviews - k * C)
                    lo, hi = shard_range(n_all, rank, world)
                    b = k & 1
                    if gather:
                        r.gather_fence(2 * b, stream.cuda_stream)        # the gathers that read this buffer pair two chunks ago
                        r.gather_fence(2 * b + 1, stream.cuda_stream)
                    if hi > lo and render:
                        # depth test off: the depth plane of every view is the constant +inf (Renderer.swift:206, the test is
                        # commented out): it is neither produced nor sent; rank 0 fills the composite's planes itself
                        r.render_device(views[base + lo:base + hi], W, H, color[b].data_ptr(), None,
                                        stream=stream.cuda_stream, sync=False)
                    if gather and world > 1:
                        gather_views(r, color[b][:hi - lo], n_all, full_c[base:base + n_all] if rank == 0 else None, slot=2 * b, stream=stream.cuda_stream)
                    elif gather:                                         # one GPU: the "composite" is a device copy
                        full_c[base:base + n_all].copy_(color[b][:n_all], non_blocking=True)
                    if gather and rank == 0:
                        full_d[base:base + n_all].fill_(inf_plane)
                if gather:
                    for slot in range(4):
                        r.gather_fence(slot, stream.cuda_stream)
                ev1.record(stream)
            r.sync()
            if gather:
                r.gather_sync()
            torch.cuda.synchronize(dev)
            return max_over_ranks(ev0.elapsed_time(ev1), device=dev)

        run(False)                                       # warm-up (policies, buffers)
        ms_render = min(run(False), run(False))
        run(True)
        ms_gather = min(run(True), run(True))
        parity = None
        if rank == 0:
            # composite against the oracle on views spread over the whole list (each rank's shard is hit)
            from oracle import oracle as O
            picks = sorted(set(int(x) for x in np.linspace(0, nviews - 1, 32)))
            parity = True
            for k in picks:
                rc, oc, od, _ = O.render(v, i, views[k], W, H)
                parity &= bool(rc == 0 and np.array_equal(full_c[k].cpu().numpy(), oc) and
                               np.array_equal(full_d[k].cpu().numpy().view(np.uint32), od.view(np.uint32)))
        # the gather alone (nothing rendered: the chunk buffers are sent as they are, so this runs AFTER the parity check - it
        # overwrites the composite with stale chunks): what NVLink + NCCL deliver into rank 0 for this message pattern
        ms_gather_only = min(run(True, render=False), run(True, render=False)) if world > 1 else 0.0
        if world > 1:
            r.comm_destroy()
    per_view = 4 * W * H                                    # colour only: the depth planes are constant and filled on rank 0
    remote = per_view * (nviews - (nviews + world - 1) // world) if world > 1 else 0       # bytes that cross NVLink into rank 0
    return {"config": workload_config(desc, W, H, ntris, nverts, False), "views": nviews, "views_per_chunk_per_rank": per_rank_chunk,
            "value": nviews / (ms_render / 1000.0), "unit": "views/s", "ms_total": ms_render,
            "frame_roofline": frame_roofline_of(ms_render / nviews * world, ntris, nverts, W, H),
            "gathered": {"value": nviews / (ms_gather / 1000.0), "unit": "views/s", "ms_total": ms_gather,
                         "gather_overhead_ms": ms_gather - ms_render, "bytes_into_rank0": remote,
                         "ingest_gbs_if_unhidden": (remote / 1e9) / max((ms_gather - ms_render) / 1000.0, 1e-9) if world > 1 and ms_gather > ms_render else None,
                         "ingest_gbs_over_whole_run": (remote / 1e9) / (ms_gather / 1000.0),
                         "gather_only_ms": ms_gather_only,
                         "gather_only_ingest_gbs": (remote / 1e9) / (ms_gather_only / 1000.0) if ms_gather_only else None,
                         "api": "sr_gather (grouped ncclSend/ncclRecv on the context's gather stream) of the colour images, chunk k gathered while "
                                "chunk k+1 renders; the depth planes (depth test off: constant +inf) are filled on rank 0, not sent",
                         "composite_parity_vs_oracle_32_views": parity}}


def run_cuda(args):
    import torch
    import torch.distributed as dist
    import software_renderer_b200 as srb
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
    if world > 1:
        dist.barrier()
    sampler = ClockSampler(local)
    sampler.start()
    per_frame, launches, t_wall = device_fps(r, xf, W, H, color, depth, stream, flush, K, Wm, frame0, dev)
    region_stats = dict(REGION_STATS)
    if world > 1:
        dist.barrier()
    clocks = sampler.stop()
    dev_ms = max_over_ranks(float(per_frame.sum()), device=dev)
    value = world * K / (dev_ms / 1000.0)

    e2e = e2e_suite(r, local, world, dev, v, i, xf, W, H, K, Wm, frame0, full=(world == 1 and not args.quick))

    line = None
    if rank == 0:
        acc = kernel_split(r, xf, W, H, color, depth, stream, flush, frame0, dev)
        stats = r.stats()
        frame_ms = dev_ms / K
        # CPU baseline: the oracle on one host core (the reference loop is sequential, Renderer.swift:222), about 10 s
        # of CPU work: calibrate on a 1/8 sample, then whole frames (or the largest 1/stride sample that fits)
        stride0 = 8 if ntris >= 500_000 else 1
        t_cal = time_oracle(v, i, xf, W, H, 1, 1, stride0, bool(args.depth_test))
        full_s = t_cal * stride0
        stride = 1
        while full_s / stride > 15.0:
            stride *= 2
        nframes = int(max(1, min(2000, (3.0 if args.quick else 10.0) / max(full_s / stride, 1e-4))))
        t_cpu = time_oracle(v, i, xf, W, H, 1, nframes, stride, bool(args.depth_test))
        cpu_value = (nframes / stride) / t_cpu
        line = {
            "metric": "frames_per_second", "value": value, "unit": "frames/s", "n_gpus": world, "steps": K, "warmup": Wm,
            "ms_per_step": frame_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32+i64", "data": "synthetic",
            "config": workload_config(desc, W, H, ntris, nverts, args.depth_test),
            "run": {"l2": "flushed between timed frames (256 MiB fill)", "parallelism": f"{world} independent view streams, one per GPU, no data-path collective"},
            "gtri_per_s": value * ntris / 1e9,
            "frame_ms_rank0": {"median": float(np.median(per_frame)), "p10": float(np.percentile(per_frame, 10)),
                               "p90": float(np.percentile(per_frame, 90))},
            "frame_roofline": frame_roofline_of(frame_ms, ntris, nverts, W, H),
            "roofline": roofline_of(acc, ntris, nverts, W, H, args.workload + ("_depth" if args.depth_test else "")),
            "kernel_ms": acc,
            "cpu_baseline": {"value": cpu_value, "unit": "frames/s", "cores": 1, "kind": "port",
                             "sample": f"{nframes} render pass(es) of every {stride}-th triangle ({ntris // stride} of {ntris} "
                                       f"tris, full clear) on 1 core, {t_cpu:.2f} s in total, each pass counted as 1/{stride} "
                                       f"frame; host has {cpu_count()} cores"},
            "e2e": e2e,
            "gpu_launches": int(launches),
            "timed_region_stats": dict(region_stats),
            "clocks": clocks,
            "wall_s_timed_region": t_wall,
            "stats_last_frame": {k: (int(x) if k != "gpu_ms" else float(x)) for k, x in stats.items()},
        }
    r.close()
    del color, depth
    # the other BASELINE configs and the depth-tested mode, 20 timed frames each (rank 0 of the one-GPU run)
    if world == 1 and not args.quick and args.workload == "soup1m" and not args.depth_test:
        wl = {}
        for key, (name, dt) in {"soup1m_depth": ("soup1m", True), "ico80k": ("ico80k", False), "cs10m4k": ("cs10m4k", False),
                                "views512": ("views512", False)}.items():
            wl[key] = workload_record(name, dt, local, dev, stream, flush)
        line["workloads"] = wl
    # configs[4], the multi-GPU configuration: sharded views, with and without the composite gather
    if not args.quick and args.workload == "soup1m" and not args.depth_test:
        c5 = cfg5_leg(local, rank, world, dev)
        if line is not None:
            line["cfg5_views512"] = c5
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if line is not None:
        OUT.emit(json.dumps(line))


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
        OUT.emit(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="cuda", choices=["cuda", "reference"])
    ap.add_argument("--workload", default="soup1m", choices=sorted(WORKLOADS))
    ap.add_argument("--depth-test", type=int, default=0)
    ap.add_argument("--split-frame", type=int, default=0, help="1: every step is ONE frame split by rows over the ranks")
    ap.add_argument("--quick", type=int, default=0, help="1: main line only (no e2e variants, no other workloads, no cfg5 leg): for kernel A/B runs")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "cuda" else args.warmup
    global OUT
    OUT = OneLineStdout()
    if args.impl == "reference":
        run_reference(args)
    elif args.split_frame:
        run_split(args)
    else:
        run_cuda(args)


if __name__ == "__main__":
    main()
