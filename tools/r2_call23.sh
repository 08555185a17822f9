#!/bin/bash
# 1 GPU: dpre recompute A/B (fixed workspace aliasing), full suite
mkdir -p gpurun_out
timeout 600 python tools/ab_step.py "dpre=0;graph=1" "dpre=1;graph=1" "dpre=0" "dpre=1" 2>&1 | grep -v Warn | tail -5
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
