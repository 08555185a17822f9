#!/bin/bash
# 1 GPU: L2 policy of the Wa rows in the BACKWARD attention kernel, now that it writes no dpre slab (r1 measured pinning there as slower)
mkdir -p gpurun_out
timeout 900 python tools/ab_step.py "l2=1,0;graph=1" "l2=1,1;graph=1" "l2=1,1;pf=0,0;graph=1" "l2=1,1;pf=0,4;graph=1" "l2=1,2;graph=1" "l2=1,1,20;graph=1" "l2=1,0" "l2=1,1" "l2=1,1;pf=0,0" 2>&1 | grep "ms/step\|rror:" | head -12
