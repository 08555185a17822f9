#!/bin/bash
# Launch list of one bench run (per-kernel gpu__time_duration under ncu; cold-cache, serialised -- shares, not absolutes).
# Run on the GPU box only after `python bench.py --steps 1 --warmup 1` exited 0 without ncu.
set -e
mkdir -p gpurun_out
timeout 300 python bench.py --steps 1 --warmup 1 > gpurun_out/bench_pre_ncu.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 1 --warmup 1 > gpurun_out/ncu_bench.log 2>&1
