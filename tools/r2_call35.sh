#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/beam_timeline.py --images 128 --beam 5 > gpurun_out/beam_timeline.txt 2>&1; grep -v Warn gpurun_out/beam_timeline.txt | head -24
