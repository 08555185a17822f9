#!/bin/bash
# 1 GPU: round-end sequence on the final state -- full GPU suite, smoke, both bench arms, beam bench (c4), beam timeline
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 400 python bench.py --impl reference > gpurun_out/bench_r2_ref_n1.json 2>/dev/null; echo "ref rc=$?"; cut -c1-120 gpurun_out/bench_r2_ref_n1.json
timeout 600 python bench.py > gpurun_out/bench_r2_n1.json 2> gpurun_out/bench_r2_n1.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.load(open('gpurun_out/bench_r2_n1.json'))
print('N=1 value', round(d['value']), 'ms', round(d['ms_per_step'], 4), 'e2e', round(d['e2e']['value']), 'cached', round(d['e2e_cached']['value']), 'frac', round(d['roofline']['frac'], 3), 'traffic', d['roofline']['traffic'], 'cpu', d['cpu_baseline']['value'], 'launches', d['gpu_launches'])
print('   beam', {k: round(v['value']) for k, v in d['beam'].items() if isinstance(v, dict) and 'value' in v} if 'error' not in d['beam'] else d['beam'])
PY
timeout 400 python tools/beam_bench.py --iters 3 --cpu-images 2 2>/dev/null | grep '^{' > gpurun_out/beam_bench_c4.jsonl; cut -c1-170 gpurun_out/beam_bench_c4.jsonl
timeout 300 python tools/beam_timeline.py --images 128 --beam 5 > gpurun_out/beam_timeline.txt 2>&1; grep -v "Warn\|_warn" gpurun_out/beam_timeline.txt | head -12
