#!/bin/bash
# 1 GPU: full suite (incl. the shape-change trainer test), bench, per-config numbers, mode switch in one process
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 300 python tools/ab_step.py "dpre=0" "dpre=1" "dpre=1;graph=1" 2>&1 | grep "ms/step\|rror:" | head
timeout 600 python bench.py > gpurun_out/bench_r2_n1.json 2> gpurun_out/bench_r2_n1.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.load(open('gpurun_out/bench_r2_n1.json'))
print('N=1 value', round(d['value']), 'ms', round(d['ms_per_step'], 4), 'e2e', round(d['e2e']['value']), 'cached', round(d['e2e_cached']['value']), 'frac', round(d['roofline']['frac'], 3), d['roofline']['kernel'], 'floor frac', round(d['roofline']['step_frac_of_hbm_floor'], 3))
print({k: (round(v.get('avg_ms', 0) * 1e3, 1), round(v.get('exposed_avg_ms', 0) * 1e3, 1)) for k, v in d['roofline']['kernels'].items() if 'avg_ms' in v})
PY
rm -f gpurun_out/config_bench_r2.jsonl
timeout 500 python tools/config_bench.py --graph 1 2>/dev/null | grep '^{' | tee -a gpurun_out/config_bench_r2.jsonl | cut -c1-100,230-330
timeout 500 python tools/config_bench.py --graph 1 --ado 1 2>/dev/null | grep '^{' | tee -a gpurun_out/config_bench_r2.jsonl | cut -c1-100,230-330
