#!/bin/bash
# 2 GPUs: data-parallel check + N=2 bench with the forked side chain (tail mode 3) as the default; full GPU suite on GPU 0
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/dp_check.py 2>&1 | grep -v "^W\|^\[W" | tail -8
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/bench_r2_n2.json 2> gpurun_out/bench_r2_n2.err; echo "bench n2 rc=$?"; tail -3 gpurun_out/bench_r2_n2.err
timeout 900 python bench.py > gpurun_out/bench_r2_n1.json 2> gpurun_out/bench_r2_n1.err; echo "bench n1 rc=$?"
python - <<'PY'
import json
for f in ('gpurun_out/bench_r2_n1.json', 'gpurun_out/bench_r2_n2.json'):
    d = json.load(open(f))
    print(f, 'value', round(d['value']), 'ms', round(d['ms_per_step'], 4), 'e2e', round(d['e2e']['value']), 'cached', round(d['e2e_cached']['value']), 'frac', round(d['roofline']['frac'], 3))
PY
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
