#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_gemm.py -m gpu -q -rf --no-header -p no:cacheprovider 2>&1 | grep -E "Greatest|Mismatched|AssertionError|FAILED|passed|failed|Error" | head -20
timeout 600 python tools/ab_step.py "wide=0" "wide=1" "wide=0;graph=1" "wide=1;graph=1" > gpurun_out/ab_wide.txt 2>&1; cat gpurun_out/ab_wide.txt
timeout 300 python - <<'PY'
import sys, torch
sys.path.insert(0, '.')
from sat_b200 import _lib
L = _lib.lib()
M, N, K = 50176, 512, 2048
g = torch.Generator(device="cuda").manual_seed(0)
A = (torch.randn(M, K, device="cuda", generator=g) * 0.5).bfloat16(); B = (torch.randn(N, K, device="cuda", generator=g) * 0.5).bfloat16()
bias = torch.randn(N, device="cuda", generator=g); C = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
st = torch.cuda.current_stream().cuda_stream
flush = torch.empty(64 << 20, device="cuda", dtype=torch.float32)
for bn, name in ((256, "128x256 tile"), (512, "256x256 tile")):
    def run():
        _lib.check(L.sat_gemm(1, M, N, K, A.data_ptr(), K, 1, B.data_ptr(), K, 1, bias.data_ptr(), 0, None, 0, C.data_ptr(), N, bn, st))
    for _ in range(3): run()
    ts = []
    for _ in range(10):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); run(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1) * 1e3)
    us = sorted(ts)[len(ts) // 2]
    ref = (A[:512].float() @ B.float().t() + bias)
    print(f"K1 {name}: {us:.1f} us = {2*M*N*K/us/1e6:.0f} TFLOP/s; max err vs fp32 on 512 rows {(C[:512].float()-ref).abs().max().item():.3e}")
PY
