#!/usr/bin/env python
"""Writes a bench workload as the flat binary tools/micro/native_loop.cu reads: dump_mesh.py <workload> <out.bin> [nmats]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
desc, W, H, v, i, xf = bench.make_workload(sys.argv[1])
n = int(sys.argv[3]) if len(sys.argv) > 3 else 64
with open(sys.argv[2], "wb") as f:
    np.array([v.shape[0], i.size], np.uint64).tofile(f)
    np.ascontiguousarray(v, np.float32).tofile(f)
    np.ascontiguousarray(i, np.int64).tofile(f)
    np.array([n], np.uint64).tofile(f)
    for k in range(n):
        np.ascontiguousarray(np.asarray(xf(k), np.float32).T).tofile(f)
print(W, H)
