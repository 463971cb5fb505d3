#!/usr/bin/env python
"""Prints the key numbers of bench JSON lines: show.py gpurun_out/bench_*.json"""
import json, sys
for p in sys.argv[1:]:
    try:
        d = json.loads(open(p).read().strip().splitlines()[-1])
    except Exception as e:
        print(p, "unreadable", e); continue
    print(f"{p}: fps={d['value']:.1f} ms={d['ms_per_step']:.4f} e2e={d['e2e']['value']:.1f} frame_frac={d.get('frame_roofline',{}).get('frac',0):.4f}")
    print("   kernels:", {k: round(v, 4) for k, v in d.get('kernel_ms', {}).items()}, "top:", d['roofline']['kernel'], f"frac={d['roofline']['frac']:.3f}")
    s = d.get('stats_last_frame', {})
    print("   stats:", {k: s[k] for k in s if k not in ('primitives', 'gpu_ms')}, "cpu=%.3f" % d['cpu_baseline']['value'], d['clocks'])
