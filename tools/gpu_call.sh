#!/bin/bash
mkdir -p gpurun_out
SR_SPREAD=2 python -m pytest tests -m gpu -x -q -k "parity or fuzz or golden or bands or boundary" > gpurun_out/c23_pytest_s2.log 2>&1; echo "SR_SPREAD=2: $(tail -1 gpurun_out/c23_pytest_s2.log)"
SR_SPREAD=3 SR_SPREAD_WAVES=4.5 python -m pytest tests -m gpu -x -q -k "parity or fuzz" > gpurun_out/c23_pytest_s3.log 2>&1; echo "SR_SPREAD=3: $(tail -1 gpurun_out/c23_pytest_s3.log)"
for rep in 1 2; do for w in soup1m ico80k; do
  for cfg in "0 1.5" "1 1.5" "2 1.5" "2 2.5" "3 4.5"; do set -- $cfg
    SR_SPREAD=$1 SR_SPREAD_WAVES=$2 python bench.py --workload $w --steps 100 --warmup 20 --quick 1 > gpurun_out/c23_${w}_s$1_w$2_r$rep.json 2>/dev/null
  done
done; done
for s in 0 2; do SR_SPREAD=$s python bench.py --workload ico80k --steps 50 --warmup 20 --quick 1 --depth-test 1 > gpurun_out/c23_icodepth_s$s.json 2>/dev/null; done
python profiles/show.py gpurun_out/c23_*.json | grep -v stats
for s in 0 2; do echo "== timeline SR_SPREAD=$s"; SR_SPREAD=$s SR_LIB=scratch/libsr_cuda_tl.so python tools/micro/py_pdl_probe.py torch 2>&1 | grep "^view" | tail -1; done
