#!/bin/bash
# 1 GPU experiments: main chain of the tail on a high-priority stream (tail=13), three-bucket update at N=1, feature upload over 2 / 4 copy streams
mkdir -p gpurun_out
timeout 600 python tools/ab_step.py "tail=3;graph=1" "tail=13;graph=1" "tail=3;buckets=3;graph=1" "tail=13;buckets=3;graph=1" "tail=3" "tail=13" 2>&1 | grep "ms/step\|rror:" | head
for n in 2 4; do
SAT_BENCH_COPY_STREAMS=$n timeout 600 python bench.py > gpurun_out/bench_copy$n.json 2> gpurun_out/bench_copy$n.err; echo "bench rc=$?"
python - <<PY
import json
d = json.load(open('gpurun_out/bench_copy$n.json'))
print('copy streams $n: value', round(d['value']), 'ms', round(d['ms_per_step'], 4), 'e2e', round(d['e2e']['value']), d['e2e']['repeats_ms_per_step'], 'h2d alone', d['e2e']['h2d_alone_ms'])
PY
done
