#!/bin/bash
mkdir -p gpurun_out
export SR_COMPACT=0 SR_LIB=scratch/libsr_cuda_tl.so
echo "pdl on, flush";   python bench.py --steps 4 --warmup 3 2>&1 >/dev/null | grep "^view" | tail -2
echo "pdl off, flush";  SR_PDL=0 python bench.py --steps 4 --warmup 3 2>&1 >/dev/null | grep "^view" | tail -2
echo "pdl on, noflush"; SR_BENCH_NOFLUSH=1 python bench.py --steps 4 --warmup 3 2>&1 >/dev/null | grep "^view" | tail -2
echo "pdl off, noflush"; SR_PDL=0 SR_BENCH_NOFLUSH=1 python bench.py --steps 4 --warmup 3 2>&1 >/dev/null | grep "^view" | tail -2
unset SR_LIB
SR_BENCH_NOFLUSH=1 python bench.py --steps 100 --warmup 20 > gpurun_out/c5_bench_noflush.json 2>/dev/null
python profiles/show.py gpurun_out/c5_bench_*.json | grep -v stats
