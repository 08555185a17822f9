#!/bin/bash
timeout 600 python tools/ab_step.py "pf=0,8;graph=1" "pf=0,4;graph=1" "pf=0,12;graph=1" "pf=0,16;graph=1" "pf=0,24;graph=1" "pf=4,8;graph=1" "pf=0,8;graph=1" 2>&1 | grep "ms/step\|rror:" | head -12
