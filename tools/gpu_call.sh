#!/bin/bash
mkdir -p gpurun_out
PIPES=sm__inst_executed_pipe_fp64.sum,sm__inst_executed_pipe_alu.sum,sm__inst_executed_pipe_fma.sum,sm__inst_executed_pipe_fmaheavy.sum,sm__inst_executed_pipe_lsu.sum,sm__inst_executed_pipe_xu.sum,sm__inst_executed.avg.per_cycle_active,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__inst_executed.sum,smsp__inst_executed.sum
# one steady frame per workload: ico80k = transform, setup, tile (3 launches); cs10m4k / views512 = transform, setup, resolve (3)
for w in ico80k cs10m4k views512; do
  python bench.py --workload $w --steps 2 --warmup 3 --quick 1 > gpurun_out/c19_plain_$w.log 2>&1 &&
  ncu --set full --metrics $PIPES --clock-control none --import-source on -k regex:'k_setup|k_tile|k_transform|k_resolve' -s 9 -c 3 -f -o gpurun_out/r02_full_$w python bench.py --workload $w --steps 2 --warmup 3 --quick 1 > gpurun_out/c19_ncu_$w.log 2>&1
done
# the shipped build's soup1m frame again (DIV32 + 6-CTA depth tile are in it now), depth off and depth LESS
python bench.py --steps 2 --warmup 3 --quick 1 > gpurun_out/c19_plain.log 2>&1 &&
ncu --set full --metrics $PIPES --clock-control none --import-source on -k regex:'k_setup|k_tile|k_transform' -s 11 -c 4 -f -o gpurun_out/r02b_full_soup1m python bench.py --steps 2 --warmup 3 --quick 1 > gpurun_out/c19_ncu.log 2>&1
python bench.py --steps 2 --warmup 3 --quick 1 --depth-test 1 > gpurun_out/c19_plain_d.log 2>&1 &&
ncu --set full --metrics $PIPES --clock-control none --import-source on -k regex:'k_setup|k_tile|k_transform' -s 9 -c 3 -f -o gpurun_out/r02b_full_soup1m_depth python bench.py --steps 2 --warmup 3 --quick 1 --depth-test 1 > gpurun_out/c19_ncu_d.log 2>&1
T0=$(date +%s); python bench.py > gpurun_out/r02_bench_1gpu.json 2> gpurun_out/r02_bench_1gpu.err; echo "default bench rc $? in $(( $(date +%s) - T0 )) s, stdout lines $(wc -l < gpurun_out/r02_bench_1gpu.json)"
T0=$(date +%s); python bench.py --impl reference > gpurun_out/r02_bench_ref.json 2> gpurun_out/r02_bench_ref.err; echo "reference arm rc $? in $(( $(date +%s) - T0 )) s"
ls -la gpurun_out | grep -E "r02b?_full"
