#!/bin/bash
timeout 300 python tools/ab_step.py "part=0" "part=2" "part=0;graph=1" "part=2;graph=1" 2>&1 | grep "ms/step\|rror:" | head
