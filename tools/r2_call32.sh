#!/bin/bash
# 1 GPU: beam search with the embedding-gate table -- parity (beam tests incl. bit-exact tokens at c4 size), then throughput
timeout 900 python -m pytest tests/test_gpu_beam.py tests/test_gpu_fullsize.py tests/test_gpu_edge_cases.py tests/test_evaluate.py -m gpu -x -q 2>&1 | tail -4
timeout 300 python tools/beam_bench.py --iters 3 2>/dev/null | grep '"fixed"' | cut -c1-150
