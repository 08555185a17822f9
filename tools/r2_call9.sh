#!/bin/bash
# ncu evidence of the round-2 state (after the plain runs exited 0 in earlier calls): full captures with source-level stall sampling of
# the attention kernels and K1 inside a real train step, of the beam-search attention / top-k kernels inside a real search, and the
# launch lists (gpu__time_duration per kernel) of one bench run and one beam search
mkdir -p gpurun_out
NCU="ncu --set full --section SourceCounters --section WarpStateStats --clock-control none --import-source on"
timeout 600 $NCU -k regex:bwd_stream_kernel -s 45 -c 1 -f -o gpurun_out/r2_attn_bwd python tools/timeline.py > gpurun_out/ncu_bwd.log 2>&1; echo "bwd rc=$?"
timeout 600 $NCU -k regex:fwd_stream_kernel -s 45 -c 1 -f -o gpurun_out/r2_attn_fwd python tools/timeline.py > gpurun_out/ncu_fwd.log 2>&1; echo "fwd rc=$?"
timeout 600 $NCU -k regex:gemm_kernel -s 185 -c 1 -f -o gpurun_out/r2_k1_gemm python tools/timeline.py > gpurun_out/ncu_k1.log 2>&1; echo "k1 rc=$?"
timeout 600 $NCU -k regex:fwd_stream_kernel -s 60 -c 1 -f -o gpurun_out/r2_beam_attn python tools/beam_timeline.py --images 128 --beam 5 > gpurun_out/ncu_beam_attn.log 2>&1; echo "beam attn rc=$?"
timeout 600 $NCU -k regex:topk_step_kernel -s 60 -c 1 -f -o gpurun_out/r2_beam_topk python tools/beam_timeline.py --images 128 --beam 5 > gpurun_out/ncu_beam_topk.log 2>&1; echo "beam topk rc=$?"
timeout 600 $NCU -k regex:tc32 -s 120 -c 1 -f -o gpurun_out/r2_beam_tc32 python tools/beam_timeline.py --images 128 --beam 5 > gpurun_out/ncu_beam_tc32.log 2>&1; echo "beam tc32 rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r2_beam_launches.csv python tools/beam_timeline.py --images 128 --beam 5 > gpurun_out/ncu_beam_list.log 2>&1; echo "beam list rc=$?"
SAT_BENCH_SKIP_SIDE=1 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r2_bench_launches.csv python bench.py --steps 2 --warmup 3 --repeats 1 > gpurun_out/ncu_bench_list.log 2>&1; echo "bench list rc=$?"
ls -la gpurun_out/*.ncu-rep gpurun_out/*.csv
# summarise on the box (the reports with sources are ~25 MB each; gpurun_out merges at most 64 MiB back)
python tools/ncu_summary.py gpurun_out/r2_ncu_train_kernels.txt attn_step_bwd=gpurun_out/r2_attn_bwd.ncu-rep attn_step_fwd=gpurun_out/r2_attn_fwd.ncu-rep k1_encoder_projection_gemm=gpurun_out/r2_k1_gemm.ncu-rep > /dev/null 2>&1
python tools/ncu_summary.py gpurun_out/r2_ncu_beam_kernels.txt beam_attn_step_fp32=gpurun_out/r2_beam_attn.ncu-rep beam_topk=gpurun_out/r2_beam_topk.ncu-rep beam_gemm_3xtf32=gpurun_out/r2_beam_tc32.ncu-rep > /dev/null 2>&1
rm -f gpurun_out/r2_attn_fwd.ncu-rep gpurun_out/r2_k1_gemm.ncu-rep gpurun_out/r2_beam_attn.ncu-rep gpurun_out/r2_beam_topk.ncu-rep gpurun_out/r2_beam_tc32.ncu-rep
head -50 gpurun_out/r2_ncu_train_kernels.txt
