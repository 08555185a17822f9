#!/bin/bash
timeout 300 python tools/ab_step.py "tail=113;graph=1" "tail=13;graph=1" "tail=113;graph=1" "tail=13;graph=1" "tail=113" "tail=13" 2>&1 | grep "ms/step\|rror:" | head
timeout 900 python -m pytest tests/test_gpu_trainer.py tests/test_gpu_decoder.py tests/test_gpu_round2.py tests/test_gpu_fullsize.py -m gpu -x -q 2>&1 | tail -3
