#!/bin/bash
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 100 --warmup 20 > gpurun_out/r02_bench_${N}gpu.json 2> gpurun_out/r02_bench_${N}gpu.err; echo "rc $? gpus $N"
python - <<PY
import json
txt = open('gpurun_out/r02_bench_${N}gpu.json').read()
print('stdout lines:', len(txt.strip().splitlines()))
d = json.loads(txt.strip().splitlines()[-1])
print('value', round(d['value']), 'ms', d['ms_per_step'], 'n', d['n_gpus'])
for k, x in d['e2e'].items():
    if k != 'api': print('  e2e', k, x if not isinstance(x, dict) else {a: b for a, b in x.items() if a != 'api'})
c = d['cfg5_views512']; g = c['gathered']
print('  cfg5 views/s', round(c['value']), 'ms', c['ms_total'], '| gathered', round(g['value']), 'ms', g['ms_total'], 'overhead', g['gather_overhead_ms'], 'GB into rank0', g['bytes_into_rank0'] / 1e9, 'gather-only ms', g['gather_only_ms'], 'GB/s', g['gather_only_ingest_gbs'], 'parity', g['composite_parity_vs_oracle_32_views'])
PY
tail -2 gpurun_out/r02_bench_${N}gpu.err
