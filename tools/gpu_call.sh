#!/bin/bash
mkdir -p gpurun_out
export SR_COMPACT=0
for m in notorch own torch; do echo "== $m"; SR_LIB=scratch/libsr_cuda_tl.so python tools/micro/py_pdl_probe.py $m 2>&1 | grep -A8 "timed frames" | grep "^view" | tail -2; done
echo "== full default bench"; T0=$(date +%s); python bench.py > gpurun_out/c8_full.json 2> gpurun_out/c8_full.err; echo "rc $? seconds $(( $(date +%s) - T0 ))"; tail -5 gpurun_out/c8_full.err
python - <<'PY'
import json
d = json.load(open('gpurun_out/c8_full.json'))
print('value', d['value'], 'ms', d['ms_per_step'])
for k, x in d['e2e'].items():
    print('  e2e', k, x if not isinstance(x, dict) else {a: b for a, b in x.items() if a != 'api'})
for k, x in d.get('workloads', {}).items():
    print('  wl', k, round(x['value'], 1), 'e2e', round(x['e2e']['value'], 1), 'parity', x['parity_vs_oracle'], 'frac', round(x['frame_roofline']['frac'], 4), x['kernel_ms'])
print('  cfg5', {k: v for k, v in d.get('cfg5_views512', {}).items() if k not in ('config',)})
PY
