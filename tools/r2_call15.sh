#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_trainer.py tests/test_gpu_round2.py tests/test_gpu_fullsize.py tests/test_gpu_gemm.py -m gpu -q -rf --no-header -p no:cacheprovider 2>&1 | grep -E "Greatest|Mismatched|AssertionError|FAILED|passed|failed|Error" | head -20
timeout 600 python tools/ab_step.py "graph=0" "graph=1" "ado=1;graph=1" > gpurun_out/ab_adam.txt 2>&1; cat gpurun_out/ab_adam.txt
