#!/bin/bash
# A/B sweep of the half-batch pipelining knobs on the bench workload: "attn_ctas,gemm_sms" pairs (0 = no cap).
mkdir -p gpurun_out
echo "split=0" ; SAT_DEC_SPLIT=0 timeout 300 python bench.py 2>/dev/null | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print(d["ms_per_step"], d["e2e"]["ms_per_step"])'
for cfg in "$@"; do
  echo "split=${MODE:-1} sms=$cfg"; SAT_DEC_SPLIT=${MODE:-1} SAT_SPLIT_SMS=$cfg timeout 300 python bench.py 2>/dev/null | python -c 'import json,sys; d=json.loads(sys.stdin.read()); k=d["roofline"]["kernels"]; print(d["ms_per_step"], d["e2e"]["ms_per_step"], k["attn_step_fwd"]["avg_ms"], k["attn_step_bwd"]["avg_ms"])'
done
