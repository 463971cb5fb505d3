#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_boundary_r2.py -m gpu -x -q 2>&1 | tail -2
for v in dm5 dm6; do SR_LIB=scratch/libsr_cuda_$v.so python -m pytest tests -m gpu -x -q -k "depth or fuzz or parity" > gpurun_out/c17_pytest_$v.log 2>&1; echo "$v: $(tail -1 gpurun_out/c17_pytest_$v.log)"; done
for rep in 1 2; do for v in base dm5 dm6; do L=""; [ $v != base ] && L=scratch/libsr_cuda_$v.so; SR_LIB=$L python bench.py --steps 50 --warmup 20 --quick 1 --depth-test 1 > gpurun_out/c17_depth_${v}_$rep.json 2>/dev/null; done; done
for v in base dm5 dm6; do L=""; [ $v != base ] && L=scratch/libsr_cuda_$v.so; SR_LIB=$L python bench.py --workload ico80k --steps 50 --warmup 20 --quick 1 --depth-test 1 > gpurun_out/c17_icodepth_$v.json 2>/dev/null; done
python profiles/show.py gpurun_out/c17_*.json | grep -v stats
