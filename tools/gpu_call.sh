#!/bin/bash
mkdir -p gpurun_out
python tools/dump_mesh.py soup1m /tmp/soup.bin 64
export SR_COMPACT=0
echo "== native loop, default lib"; scratch/native_loop software_renderer_b200/libsr_cuda.so /tmp/soup.bin 1920 1080 100
echo "== native loop, PDL off"; SR_PDL=0 scratch/native_loop software_renderer_b200/libsr_cuda.so /tmp/soup.bin 1920 1080 100
echo "== native loop, TL lib"; scratch/native_loop scratch/libsr_cuda_tl.so /tmp/soup.bin 1920 1080 4 2>&1 | grep -E "^view|native" | tail -3
echo "== native loop, TL lib, PDL off"; SR_PDL=0 scratch/native_loop scratch/libsr_cuda_tl.so /tmp/soup.bin 1920 1080 4 2>&1 | grep -E "^view|native" | tail -3
echo "== python, TL lib"; SR_LIB=scratch/libsr_cuda_tl.so python bench.py --steps 4 --warmup 3 2>&1 >/dev/null | grep "^view" | tail -2
