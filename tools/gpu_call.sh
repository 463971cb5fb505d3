#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/c13_pytest.log 2>&1; tail -3 gpurun_out/c13_pytest.log
for v in DIV32 VEC_WB; do SR_LIB=scratch/libsr_cuda_$v.so python -m pytest tests -m gpu -x -q -k "parity or fuzz or golden or bands" > gpurun_out/c13_pytest_$v.log 2>&1; echo "$v: $(tail -1 gpurun_out/c13_pytest_$v.log)"; done
for w in soup1m ico80k cs10m4k; do
  python bench.py --workload $w --steps 100 --warmup 20 --quick 1 > gpurun_out/c13_ab_base_$w.json 2>/dev/null
  for v in DIV32 VEC_WB; do SR_LIB=scratch/libsr_cuda_$v.so python bench.py --workload $w --steps 100 --warmup 20 --quick 1 > gpurun_out/c13_ab_${v}_$w.json 2>/dev/null; done
done
python bench.py --workload soup1m --steps 100 --warmup 20 --quick 1 > gpurun_out/c13_ab_base2_soup1m.json 2>/dev/null
python profiles/show.py gpurun_out/c13_ab_*.json | grep -v stats
PIPES=sm__inst_executed_pipe_fp64.sum,sm__inst_executed_pipe_alu.sum,sm__inst_executed_pipe_fma.sum,sm__inst_executed_pipe_fmaheavy.sum,sm__inst_executed_pipe_lsu.sum,sm__inst_executed_pipe_xu.sum,sm__inst_executed.avg.per_cycle_active,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__inst_executed.sum,smsp__inst_executed.sum
python bench.py --steps 2 --warmup 3 --quick 1 > gpurun_out/c13_plain.log 2>&1 &&
ncu --set full --metrics $PIPES --clock-control none --import-source on -k regex:'k_setup|k_tile|k_transform' -s 11 -c 4 -f -o gpurun_out/r02_full_soup1m python bench.py --steps 2 --warmup 3 --quick 1 > gpurun_out/c13_ncu.log 2>&1
python bench.py --steps 2 --warmup 3 --quick 1 > gpurun_out/c13_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02_launches_soup1m.csv python bench.py --steps 2 --warmup 3 --quick 1 > /dev/null 2>&1
python bench.py --steps 2 --warmup 3 --quick 1 --depth-test 1 > gpurun_out/c13_plain_d.log 2>&1 &&
ncu --set full --metrics $PIPES --clock-control none --import-source on -k regex:'k_setup|k_tile|k_transform' -s 9 -c 3 -f -o gpurun_out/r02_full_soup1m_depth python bench.py --steps 2 --warmup 3 --quick 1 --depth-test 1 > gpurun_out/c13_ncu_d.log 2>&1
ls -la gpurun_out | grep r02_
