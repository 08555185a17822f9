#!/bin/bash
# 1 GPU: backward tail with the side chain forked onto a second stream (sat_set_tail_overlap(3)) -- A/B, parity, timeline
mkdir -p gpurun_out
timeout 600 python tools/ab_step.py "tail=1;graph=1" "tail=3;graph=1" "tail=1" "tail=3" "tail=1;graph=1" "tail=3;graph=1" 2>&1 | grep -v Warn | tail -8
timeout 900 python -c "
import sys; sys.path.insert(0, '.')
from sat_b200 import _lib
_lib.lib().sat_set_tail_overlap(3)
import pytest
sys.exit(pytest.main(['tests/test_gpu_trainer.py', 'tests/test_gpu_decoder.py', 'tests/test_gpu_round2.py', 'tests/test_gpu_fullsize.py', 'tests/test_gpu_edge_cases.py', '-m', 'gpu', '-x', '-q']))
" 2>&1 | tail -4
SAT_TAIL=3 timeout 300 python tools/timeline.py > gpurun_out/timeline_tail3.txt 2>&1; tail -30 gpurun_out/timeline_tail3.txt | head -5
