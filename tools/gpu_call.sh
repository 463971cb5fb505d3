#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/c18_pytest.log 2>&1; tail -2 gpurun_out/c18_pytest.log
for v in dm8 sd4 sd6 sd8; do SR_LIB=scratch/libsr_cuda_$v.so python -m pytest tests -m gpu -x -q -k "depth or fuzz" > gpurun_out/c18_pytest_$v.log 2>&1; echo "$v: $(tail -1 gpurun_out/c18_pytest_$v.log)"; done
for rep in 1 2; do for v in base dm8 sd4 sd6 sd8; do L=""; [ $v != base ] && L=scratch/libsr_cuda_$v.so; SR_LIB=$L python bench.py --steps 50 --warmup 20 --quick 1 --depth-test 1 > gpurun_out/c18_depth_${v}_$rep.json 2>/dev/null; done; done
for v in base dm8 sd6; do L=""; [ $v != base ] && L=scratch/libsr_cuda_$v.so; SR_LIB=$L python bench.py --workload ico80k --steps 50 --warmup 20 --quick 1 --depth-test 1 > gpurun_out/c18_icodepth_$v.json 2>/dev/null; done
python profiles/show.py gpurun_out/c18_*.json | grep -v stats
