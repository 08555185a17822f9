#!/bin/bash
mkdir -p gpurun_out
rm -f gpurun_out/fullsize_parity.jsonl
timeout 900 python -m pytest tests/test_gpu_round2.py tests/test_gpu_fullsize.py tests/test_gpu_beam.py tests/test_gpu_attention.py -m gpu -q -rf --no-header -p no:cacheprovider > gpurun_out/pytest_gpu2.log 2>&1; echo "pytest rc=$?"
tail -12 gpurun_out/pytest_gpu2.log
timeout 600 python tools/ab_step.py "pf=0,0" "pf=16,16" "pf=32,32" "pf=64,64" "pf=128,128" "pf=32,0" "pf=0,32" "part=2" "pf=32,32;part=2" "pf=64,64;part=2" "graph=1" "pf=32,32;graph=1" "ado=1" > gpurun_out/ab_step.txt 2>&1; echo "ab rc=$?"
cat gpurun_out/ab_step.txt
timeout 300 python tools/beam_timeline.py > gpurun_out/beam_timeline2.txt 2>&1; grep "^# dtype" gpurun_out/beam_timeline2.txt
timeout 300 python tools/timeline.py --full > gpurun_out/timeline_full.txt 2>&1; tail -30 gpurun_out/timeline_full.txt
