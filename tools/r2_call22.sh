#!/bin/bash
# 1 GPU: sum_t dpre rebuilt from the score gradients (sat_set_dpre_recompute) -- parity first, then A/B and a timeline
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_trainer.py tests/test_gpu_decoder.py tests/test_gpu_round2.py tests/test_gpu_fullsize.py tests/test_gpu_edge_cases.py tests/test_gpu_attention.py -m gpu -x -q 2>&1 | tail -12
timeout 600 python tools/ab_step.py "dpre=0;graph=1" "dpre=1;graph=1" "dpre=0" "dpre=1" 2>&1 | grep -v Warn | tail -5
timeout 300 python tools/timeline.py --full > gpurun_out/timeline_dpre.txt 2>&1; grep "step span" gpurun_out/timeline_dpre.txt
