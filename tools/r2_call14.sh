#!/bin/bash
# 8 GPUs, exactly the driver's launch line: reference arm first (rank 0 only), then ours
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531 bench.py --impl reference --gpus 8 --steps 3 --warmup 1 2>/dev/null | cut -c1-200
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29532 bench.py --gpus 8 --steps 20 --warmup 3 > gpurun_out/bench_r2_n8.json 2> gpurun_out/bench_r2_n8.err; echo "bench n8 rc=$?"; grep -v "^\[W\|^W1\|_warn\|UserWarning\|NCCL version" gpurun_out/bench_r2_n8.err | tail -8
python - <<'PY'
import json
d = json.load(open('gpurun_out/bench_r2_n8.json'))
print('N=8 value', round(d['value']), 'ms', round(d['ms_per_step'], 4), 'e2e', round(d['e2e']['value']), round(d['e2e']['ms_per_step'], 3), 'cached', round(d['e2e_cached']['value']), round(d['e2e_cached']['ms_per_step'], 3), 'h2d GB/s', round(d['e2e']['h2d_alone_gbs'], 1), d['config']['host_numa'])
print('   beam', {k: round(v['value']) for k, v in d['beam'].items() if isinstance(v, dict) and 'value' in v} if 'error' not in d['beam'] else d['beam'])
print('   launch:', d['config']['launch'][:60], '| exchange:', d['config']['grad_exchange'])
PY
