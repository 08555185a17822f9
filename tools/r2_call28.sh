#!/bin/bash
# 1 GPU: cp.async-pipelined dpre rebuild (default) vs the plain kernel (dpre=2): parity, A/B, isolated duration
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_trainer.py tests/test_gpu_decoder.py tests/test_gpu_round2.py tests/test_gpu_fullsize.py tests/test_gpu_edge_cases.py -m gpu -x -q 2>&1 | tail -3
timeout 600 python tools/ab_step.py "dpre=2;graph=1" "dpre=1;graph=1" "dpre=2;graph=1" "dpre=1;graph=1" 2>&1 | grep "ms/step\|rror:" | head
SAT_TAIL=0 timeout 300 python tools/timeline.py --full > gpurun_out/timeline_tail0.txt 2>&1; grep "dpre_sum" gpurun_out/timeline_tail0.txt | head -2
