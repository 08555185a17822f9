#!/bin/bash
# round-2 GPU call 1: all GPU tests (old + new), kineto timelines of the train step and of beam search, ncu full capture of the
# backward attention kernel with source-level stall sampling, ncu launch list of one beam-search call
mkdir -p gpurun_out
rm -f gpurun_out/fullsize_parity.jsonl
timeout 1200 python -m pytest tests -m gpu -q -rf --no-header -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/pytest_gpu.log
timeout 300 python tools/beam_timeline.py > gpurun_out/beam_timeline.txt 2>&1; echo "beam_timeline rc=$?"
timeout 300 python tools/timeline.py --full > gpurun_out/timeline_full.txt 2>&1; echo "timeline rc=$?"
timeout 600 ncu --set full --section SourceCounters --section WarpStateStats --clock-control none --import-source on -k regex:bwd_stream_kernel -s 45 -c 1 -f -o gpurun_out/attn_bwd_r2 \
    python tools/timeline.py > gpurun_out/ncu_full_bwd.log 2>&1; echo "ncu bwd rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/beam_launches.csv \
    python tools/beam_timeline.py --images 128 --beam 5 > gpurun_out/ncu_beam.log 2>&1; echo "ncu beam rc=$?"
ls -la gpurun_out/
