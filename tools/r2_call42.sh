#!/bin/bash
# N GPUs (NG), the driver's launch line, final state: reference arm (rank 0 only), then ours
NG=${NG:-8}
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29531 bench.py --impl reference --gpus $NG --steps 3 --warmup 1 2>/dev/null | cut -c1-160
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29532 bench.py --gpus $NG --steps 20 --warmup 3 > gpurun_out/bench_r2_n$NG.json 2> gpurun_out/bench_r2_n$NG.err; echo "bench n$NG rc=$?"; grep -v "^\[W\|^W1\|_warn\|UserWarning\|NCCL version\|^\*\|OMP_NUM\|^$" gpurun_out/bench_r2_n$NG.err | tail -5
python - <<PY
import json
d = json.load(open('gpurun_out/bench_r2_n$NG.json'))
print('N=$NG value', round(d['value']), 'ms', round(d['ms_per_step'], 4), d['config'].get('repeats_ms_per_step'), 'e2e', round(d['e2e']['value']), 'cached', round(d['e2e_cached']['value']), '|', d['config']['grad_exchange'][:70])
print('   beam', {k: round(v['value']) for k, v in d['beam'].items() if isinstance(v, dict) and 'value' in v} if 'error' not in d['beam'] else d['beam'])
PY
