#!/bin/bash
mkdir -p gpurun_out
for v in "dpre=0;graph=1" "dpre=1;graph=1" "dpre=0" "dpre=1"; do
  timeout 300 python tools/ab_step.py "$v" > gpurun_out/ab_one.log 2>&1; echo "rc=$? $v"; grep -v Warn gpurun_out/ab_one.log | grep "ms/step\|Error\|error" | head -3
done
timeout 300 python tools/ab_step.py "dpre=0" "dpre=1" > gpurun_out/ab_two.log 2>&1; echo "rc=$?"; grep "ms/step\|Error\|rror:" gpurun_out/ab_two.log | head; grep -n "File \"/root/repo\|File \"tools" gpurun_out/ab_two.log | tail -5
