#!/bin/bash
mkdir -p gpurun_out
nvidia-smi -L | wc -l
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 8 --steps 100 --warmup 20 > gpurun_out/r02_bench_8gpu.json 2> gpurun_out/r02_bench_8gpu.err; echo "rc $?"
python - <<'PY'
import json
txt = open('gpurun_out/r02_bench_8gpu.json').read()
print('stdout lines:', len(txt.strip().splitlines()))
d = json.loads([l for l in txt.splitlines() if l.startswith('{')][-1])
print('value', round(d['value']), 'ms', d['ms_per_step'], 'n', d['n_gpus'])
for k, x in d['e2e'].items():
    if k != 'api': print('  e2e', k, x if not isinstance(x, dict) else {a: b for a, b in x.items() if a != 'api'})
c = d['cfg5_views512']; g = c['gathered']
print('  cfg5 views/s', round(c['value']), 'ms', c['ms_total'], '| gathered', round(g['value']), 'ms', g['ms_total'], 'overhead', g['gather_overhead_ms'], 'GB into rank0', g['bytes_into_rank0'] / 1e9, 'GB/s whole run', g['ingest_gbs_over_whole_run'], 'parity', g['composite_parity_vs_oracle_32_views'])
PY
python tools/pcie_probe.py --sweep 1,2,4,8 > gpurun_out/r02_pcie_probe.jsonl 2>/dev/null
python - <<'PY'
import json
for l in open('gpurun_out/r02_pcie_probe.jsonl'):
    d = json.loads(l)
    if 'pinned' in d: print(d['n_gpus'], d['block_bytes'], 'D2H pinned agg', d['pinned']['gbs_aggregate'], 'per', d['pinned']['gbs_per_gpu'][:2], '| pageable agg', d['pageable']['gbs_aggregate'], '| H2D agg', d['h2d_pinned']['gbs_aggregate'])
PY
tail -2 gpurun_out/r02_bench_8gpu.err
