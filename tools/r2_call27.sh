#!/bin/bash
# 1 GPU: final-state evidence -- GPU suite, bench (plain), then ncu: full captures (stall sampling) of the attention kernels, the dpre
# rebuild and K1 inside a real train step, and the launch list of one bench run
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 600 python bench.py > gpurun_out/bench_r2_n1.json 2> gpurun_out/bench_r2_n1.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.load(open('gpurun_out/bench_r2_n1.json'))
print('N=1 value', round(d['value']), 'ms', round(d['ms_per_step'], 4), 'e2e', round(d['e2e']['value']), 'cached', round(d['e2e_cached']['value']), 'frac', round(d['roofline']['frac'], 3), 'launches/step', d['gpu_launches_per_step'])
PY
SAT_TAIL=0 timeout 300 python tools/timeline.py --full > gpurun_out/timeline_tail0.txt 2>&1; grep "dpre_sum" gpurun_out/timeline_tail0.txt | head -3
NCU="ncu --set full --section SourceCounters --section WarpStateStats --clock-control none --import-source on"
timeout 600 $NCU -k regex:bwd_stream_kernel -s 45 -c 1 -f -o gpurun_out/r2_attn_bwd python tools/timeline.py > gpurun_out/ncu_bwd.log 2>&1; echo "bwd rc=$?"
timeout 600 $NCU -k regex:fwd_stream_kernel -s 45 -c 1 -f -o gpurun_out/r2_attn_fwd python tools/timeline.py > gpurun_out/ncu_fwd.log 2>&1; echo "fwd rc=$?"
timeout 600 $NCU -k regex:dpre_sum_kernel -s 2 -c 1 -f -o gpurun_out/r2_dpre_sum python tools/timeline.py > gpurun_out/ncu_dpre.log 2>&1; echo "dpre rc=$?"
timeout 600 $NCU -k regex:gemm2m_kernel -s 2 -c 1 -f -o gpurun_out/r2_k1_gemm python tools/timeline.py > gpurun_out/ncu_k1.log 2>&1; echo "k1 rc=$?"
SAT_BENCH_SKIP_SIDE=1 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r2_bench_launches.csv python bench.py --steps 2 --warmup 3 --repeats 1 > gpurun_out/ncu_bench_list.log 2>&1; echo "bench list rc=$?"
python tools/ncu_summary.py gpurun_out/r2_ncu_train_kernels.txt attn_step_bwd=gpurun_out/r2_attn_bwd.ncu-rep attn_step_fwd=gpurun_out/r2_attn_fwd.ncu-rep dpre_sum=gpurun_out/r2_dpre_sum.ncu-rep k1_encoder_projection_gemm=gpurun_out/r2_k1_gemm.ncu-rep > /dev/null 2>&1
rm -f gpurun_out/*.ncu-rep
grep -n "^## \|gpu__time_duration\|dram__bytes_read.sum \|dram__bytes_write.sum \|pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed" gpurun_out/r2_ncu_train_kernels.txt | head -40
