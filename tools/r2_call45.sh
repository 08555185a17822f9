#!/bin/bash
timeout 600 python tools/ab_step.py "l2=1,0;pf=0,8;graph=1" "l2=1,1;pf=0,1008;graph=1" "l2=1,0;pf=0,1008;graph=1" "l2=1,2;pf=0,1008;graph=1" "l2=1,1;pf=0,1016;graph=1" "l2=1,1;pf=1008,1008;graph=1" "l2=1,0;pf=0,8" "l2=1,1;pf=0,1008" 2>&1 | grep "ms/step\|rror:" | head -12
