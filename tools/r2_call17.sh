#!/bin/bash
# 1 GPU: the round-end sequence -- full GPU suite, smoke, both bench arms at the driver's defaults, per-config numbers
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 400 python bench.py --impl reference > gpurun_out/bench_r2_ref_n1.json 2>/dev/null; echo "ref rc=$?"; cut -c1-300 gpurun_out/bench_r2_ref_n1.json
timeout 600 python bench.py > gpurun_out/bench_r2_n1.json 2> gpurun_out/bench_r2_n1.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.load(open('gpurun_out/bench_r2_n1.json'))
print('N=1 value', round(d['value']), 'ms', round(d['ms_per_step'], 4), 'e2e', round(d['e2e']['value']), 'cached', round(d['e2e_cached']['value']), 'frac', round(d['roofline']['frac'], 3), 'cpu', d['cpu_baseline']['value'], 'launches', d['gpu_launches'])
print('   beam', {k: round(v['value']) for k, v in d['beam'].items() if isinstance(v, dict) and 'value' in v} if 'error' not in d['beam'] else d['beam'])
print('   kernels', {k: round(v.get('avg_ms', 0) * 1e3, 1) for k, v in d['roofline']['kernels'].items() if 'avg_ms' in v})
PY
timeout 500 python tools/config_bench.py --graph 1 2>/dev/null | grep '^{' | tee -a gpurun_out/config_bench_r2.jsonl | cut -c1-260
timeout 500 python tools/config_bench.py --graph 1 --ado 1 2>/dev/null | grep '^{' | tee -a gpurun_out/config_bench_r2.jsonl | cut -c1-260
