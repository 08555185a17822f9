#!/bin/bash
# 2 GPUs: data-parallel check, then bench at N=2 exactly as the driver launches it; N=1 bench and reference arm on the same box
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/dp_check.py 2>&1 | grep -v "^W\|^\[W" | tail -8
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/bench_r2_n2.json 2> gpurun_out/bench_r2_n2.err; echo "bench n2 rc=$?"; tail -5 gpurun_out/bench_r2_n2.err
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 3 > gpurun_out/bench_r2_n1.json 2> gpurun_out/bench_r2_n1.err; echo "bench n1 rc=$?"; tail -3 gpurun_out/bench_r2_n1.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --impl reference --gpus 2 --steps 3 --warmup 1 2>/dev/null | cut -c1-300
python - <<'PY'
import json
for f in ('gpurun_out/bench_r2_n1.json', 'gpurun_out/bench_r2_n2.json'):
    try:
        d = json.load(open(f))
    except Exception as e:
        print(f, 'unreadable', e); continue
    print(f, 'value', round(d['value']), 'ms', round(d['ms_per_step'], 4), 'launches/step', d['gpu_launches_per_step'], '|', d['config']['launch'][:40])
    print('   e2e', round(d['e2e']['value']), round(d['e2e']['ms_per_step'], 3), 'cached', round(d['e2e_cached']['value']), round(d['e2e_cached']['ms_per_step'], 3), 'h2d GB/s', round(d['e2e']['h2d_alone_gbs'], 1), d['config']['host_numa'])
    r = d['roofline']
    print('   roofline', r['kernel'], round(r['frac'], 3), 'step frac of floor', round(r['step_frac_of_hbm_floor'], 3), {k: (round(v.get('avg_ms', 0), 4), round(v.get('exposed_avg_ms', 0), 4)) for k, v in r['kernels'].items() if 'avg_ms' in v})
    print('   cpu', d['cpu_baseline'] and (round(d['cpu_baseline']['value'], 1), d['cpu_baseline']['kind']), 'clocks', d['clocks'] and (d['clocks']['sm_mhz'], d['clocks']['reasons']))
    print('   beam', {k: round(v['value']) for k, v in d['beam'].items() if isinstance(v, dict) and 'value' in v} if 'error' not in d['beam'] else d['beam'])
PY
