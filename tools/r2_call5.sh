#!/bin/bash
mkdir -p gpurun_out
echo "== alone, fresh process"; timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q --no-header -p no:cacheprovider 2>&1 | tail -15
echo "== full suite"; rm -f gpurun_out/fullsize_parity.jsonl
timeout 1500 python -m pytest tests -m gpu -q -rf --no-header -p no:cacheprovider > gpurun_out/pytest_gpu_r2b.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu_r2b.log
