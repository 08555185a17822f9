#!/bin/bash
# ncu evidence of the current state (run on the GPU box AFTER `python bench.py` has exited 0 without ncu):
#   1. launch list of one bench run (gpu__time_duration per kernel; cold cache, serialised: shares, not absolutes)
#   2. --set full captures of the attention step kernels and of K1 (encoder projection GEMM) inside a real train step
set -e
mkdir -p gpurun_out
timeout 300 python bench.py --steps 2 --warmup 3 > gpurun_out/bench_pre_ncu.json 2> gpurun_out/bench_pre_ncu.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 3 > gpurun_out/ncu_bench.log 2>&1 || true
# the workload of the full captures: tools/timeline.py runs 6 fused train steps of the bench workload
timeout 600 ncu --set full --clock-control none --import-source on -k regex:bwd_stream_kernel -s 45 -c 1 -f -o gpurun_out/attn_bwd \
    python tools/timeline.py > gpurun_out/ncu_full_bwd.log 2>&1 || true
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fwd_stream_kernel -s 45 -c 1 -f -o gpurun_out/attn_fwd \
    python tools/timeline.py > gpurun_out/ncu_full_fwd.log 2>&1 || true
# K1 is the 2nd tcgen05 GEMM of every step (92 GEMM launches per step): skip 2 steps + 1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_kernel -s 185 -c 1 -f -o gpurun_out/k1_gemm \
    python tools/timeline.py > gpurun_out/ncu_full_k1.log 2>&1 || true
ls -la gpurun_out/*.ncu-rep
