#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_trainer.py tests/test_gpu_beam.py tests/test_gpu_decoder.py tests/test_gpu_fullsize.py -m gpu -q -rf --no-header -p no:cacheprovider 2>&1 | grep -E "Greatest|Mismatched|AssertionError: \(|FAILED|passed|failed" | head -30
timeout 300 python - <<'PY'
# accuracy + speed of the fp32 GEMM paths on the beam-search LSTM shape (M=640, N=2048, K=1536) and a long-K case
import sys, torch
sys.path.insert(0, '.')
from sat_b200 import _lib
L = _lib.lib()
g = torch.Generator(device="cuda").manual_seed(0)
for (M, N, K) in ((640, 2048, 1536), (256, 2048, 2560), (640, 2536, 512)):
    A = torch.randn(M, K, device="cuda", generator=g) * 0.5
    B = torch.randn(N, K, device="cuda", generator=g) * 0.5
    ref = (A.double() @ B.double().t())
    out = {}
    for mode in (0, 1):
        L.sat_set_fp32_gemm_mode(mode)
        C = torch.empty(M, N, device="cuda")
        st = torch.cuda.current_stream().cuda_stream
        def run():
            _lib.check(L.sat_gemm(0, M, N, K, A.data_ptr(), K, 1, B.data_ptr(), K, 1, None, 0, C.data_ptr(), N, None, 0, 0, st))
        for _ in range(3): run()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20): run()
        e1.record(); torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / 20 * 1e3
        err = (C.double() - ref).abs().max().item()
        out[mode] = (us, err)
    t = (A @ B.t()).double()
    print(f"M={M} N={N} K={K}: simt {out[0][0]:.1f} us max|err| {out[0][1]:.2e};  3xTF32 {out[1][0]:.1f} us max|err| {out[1][1]:.2e};  torch fp32 max|err| {(t-ref).abs().max().item():.2e}; |C|max {ref.abs().max().item():.1f}  ({2*M*N*K/out[1][0]/1e6:.1f} TFLOP/s fp32-equivalent)")
L.sat_set_fp32_gemm_mode(1)
PY
