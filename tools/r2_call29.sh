#!/bin/bash
# 8 GPUs, the driver's launch line: reference arm (rank 0 only), ours with the default two-bucket exchange, then the three-bucket variant
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531 bench.py --impl reference --gpus 8 --steps 3 --warmup 1 2>/dev/null | cut -c1-200
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29532 bench.py --gpus 8 --steps 20 --warmup 3 > gpurun_out/bench_r2_n8.json 2> gpurun_out/bench_r2_n8.err; echo "bench n8 rc=$?"; grep -v "^\[W\|^W1\|_warn\|UserWarning\|NCCL version" gpurun_out/bench_r2_n8.err | tail -5
SAT_BUCKETS=3 SAT_BENCH_SKIP_SIDE=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 8 --steps 20 --warmup 3 > gpurun_out/bench_r2_n8_b3.json 2> gpurun_out/bench_r2_n8_b3.err; echo "bench n8 b3 rc=$?"
python - <<'PY'
import json
for f in ('gpurun_out/bench_r2_n8.json', 'gpurun_out/bench_r2_n8_b3.json'):
    try:
        d = json.load(open(f))
    except Exception as e:
        print(f, 'unreadable', e); continue
    print(f, 'value', round(d['value']), 'ms', round(d['ms_per_step'], 4), d['config'].get('repeats_ms_per_step'), 'e2e', round(d['e2e']['value']), 'cached', round(d['e2e_cached']['value']), '|', d['config']['grad_exchange'][:50])
    print('   beam', {k: round(v['value']) for k, v in d['beam'].items() if isinstance(v, dict) and 'value' in v} if 'error' not in d['beam'] else d['beam'])
PY
