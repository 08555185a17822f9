#!/bin/bash
# 1 GPU: fp32 beam attention with tanh from separate exponentials -- attention / beam parity, then throughput
timeout 900 python -m pytest tests/test_gpu_attention.py tests/test_gpu_beam.py tests/test_gpu_fullsize.py tests/test_gpu_edge_cases.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -4
timeout 300 python tools/beam_bench.py --iters 3 2>/dev/null | grep '"fixed"' | cut -c1-150
