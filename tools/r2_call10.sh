#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_beam.py tests/test_gpu_attention.py tests/test_gpu_decoder.py tests/test_gpu_fullsize.py tests/test_gpu_round2.py tests/test_zz_no_alloc.py -m gpu -q -rf --no-header -p no:cacheprovider 2>&1 | grep -E "Greatest|Mismatched|AssertionError|FAILED|passed|failed|Error" | head -30
timeout 300 python tools/beam_timeline.py > gpurun_out/beam_timeline_r2b.txt 2>&1; grep -A9 "^# dtype" gpurun_out/beam_timeline_r2b.txt
timeout 300 python tools/beam_bench.py --images 128 --iters 3 > gpurun_out/beam_bench_r2.jsonl 2>gpurun_out/beam_bench_r2.err; cut -c1-220 gpurun_out/beam_bench_r2.jsonl
timeout 300 python tools/timeline.py --full > gpurun_out/timeline_r2.txt 2>&1; tail -45 gpurun_out/timeline_r2.txt
