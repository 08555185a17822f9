#!/bin/bash
mkdir -p gpurun_out
rm -f gpurun_out/fullsize_parity.jsonl
timeout 1500 python -m pytest tests -m gpu -q -rf --no-header -p no:cacheprovider > gpurun_out/pytest_gpu_r2c.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/pytest_gpu_r2c.log
timeout 600 python tools/ab_step.py "pf=0,0" "pf=0,8" "pf=0,16" "pf=8,8" "pf=16,16" "pf=0,8;part=2" "graph=1;pf=0,8" "ado=1;pf=0,8" > gpurun_out/ab_step_r2.txt 2>&1; echo "ab rc=$?"; cat gpurun_out/ab_step_r2.txt
timeout 300 python tools/beam_timeline.py > gpurun_out/beam_timeline_r2.txt 2>&1; grep -A8 "^# dtype" gpurun_out/beam_timeline_r2.txt
