#!/bin/bash
# 1 GPU: beam search at c4 with whole-item (automatic) vs equal-slot partition of the attention stream (128 images on 148 SMs)
for m in 0 2; do echo "part=$m"; SAT_ATTN_PART=$m timeout 300 python tools/beam_bench.py --iters 3 2>/dev/null | grep '"fixed"' | cut -c1-210; done
