#!/bin/bash
# 8 GPUs: gradient exchange as ONE all-reduce after the pass (SAT_OVERLAP=0) vs the default two buckets, same box, device-resident value only
mkdir -p gpurun_out
NG=${NG:-8}
for v in 1 0 1 0; do
SAT_OVERLAP=$v SAT_BENCH_SKIP_SIDE=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 2954$v bench.py --gpus $NG --steps 20 --warmup 3 > gpurun_out/bench_n8_ov$v.json 2> gpurun_out/bench_n8_ov$v.err; echo "overlap=$v rc=$?"
python - <<PY
import json
d = json.load(open('gpurun_out/bench_n8_ov$v.json'))
print('  value', round(d['value']), 'ms', round(d['ms_per_step'], 4), d['config'].get('repeats_ms_per_step'), '|', d['config']['grad_exchange'][:60])
PY
done
