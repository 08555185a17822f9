#!/bin/bash
# 1 GPU: full GPU suite + the bench line (after the e2e staging change)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
timeout 600 python bench.py > gpurun_out/bench_r2_n1.json 2> gpurun_out/bench_r2_n1.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_r2_n1.err
python - <<'PY'
import json
d = json.load(open('gpurun_out/bench_r2_n1.json'))
print('N=1 value', round(d['value']), 'ms', round(d['ms_per_step'], 4), 'e2e', round(d['e2e']['value']), d['e2e']['repeats_ms_per_step'], 'cached', round(d['e2e_cached']['value']), 'frac', round(d['roofline']['frac'], 3), 'cpu', d['cpu_baseline']['value'], 'launches', d['gpu_launches'])
PY
