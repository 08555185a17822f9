"""GPU: sat_gemm (fp32 SIMT kernel and bf16 tcgen05 kernel) vs torch matmul, all operand layouts, ragged sizes."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _run(dt, M, N, K, a_k, b_k, act=0, bias=True, bn=0, seed=0):
    from sat_b200 import _lib
    L = _lib.lib()
    g = torch.Generator(device="cuda").manual_seed(seed)
    A = torch.randn(M, K, device="cuda", generator=g) * 0.5
    B = torch.randn(N, K, device="cuda", generator=g) * 0.5
    bvec = torch.randn(N, device="cuda", generator=g) if bias else None
    Ad, Bd = A.to(dt), B.to(dt)
    ref = Ad.float() @ Bd.float().t()
    if bias:
        ref = ref + bvec
    if act == 1:
        ref = torch.tanh(ref)
    elif act == 2:
        ref = torch.relu(ref)
    # pad leading dims to multiples of 8 (TMA pitch requirement), keep logical sizes ragged
    def store(X, kmajor):
        Xs = X if kmajor else X.t()
        r, c = Xs.shape
        ld = (c + 7) // 8 * 8
        buf = torch.zeros(r, ld, device="cuda", dtype=dt)
        buf[:, :c] = Xs
        return buf, ld
    As, lda = store(Ad, a_k)
    Bs, ldb = store(Bd, b_k)
    ldc = (N + 7) // 8 * 8
    Cf = torch.full((M, ldc), float("nan"), device="cuda")
    Ct = torch.zeros(M, ldc, device="cuda", dtype=dt)
    rc = L.sat_gemm(_lib.dtype_code(dt), M, N, K, As.data_ptr(), lda, int(a_k), Bs.data_ptr(), ldb, int(b_k),
                    None if bvec is None else bvec.data_ptr(), act, Cf.data_ptr(), ldc, Ct.data_ptr(), ldc, bn,
                    torch.cuda.current_stream().cuda_stream)
    _lib.check(rc, "sat_gemm")
    torch.cuda.synchronize()
    tol = dict(rtol=1e-4, atol=1e-4) if dt == torch.float32 else dict(rtol=2e-3, atol=2e-3 * (K ** 0.5) * 0.25 + 1e-3)
    torch.testing.assert_close(Cf[:, :N], ref, **tol)
    torch.testing.assert_close(Ct[:, :N].float(), ref, rtol=1e-2, atol=max(tol["atol"], 2e-2) if dt != torch.float32 else 1e-4)


SHAPES = [(32, 64, 64), (128, 128, 128), (200, 72, 136), (256, 2048, 2560), (77, 2531, 512), (512, 100, 1000), (1, 8, 8)]


@pytest.mark.parametrize("M,N,K", SHAPES)
@pytest.mark.parametrize("a_k,b_k", [(True, True), (True, False), (False, True), (False, False)])
def test_gemm_fp32(M, N, K, a_k, b_k):
    _run(torch.float32, M, N, K, a_k, b_k)


@pytest.mark.parametrize("M,N,K", SHAPES)
@pytest.mark.parametrize("a_k,b_k", [(True, True), (True, False), (False, True), (False, False)])
def test_gemm_bf16_tcgen05(M, N, K, a_k, b_k):
    _run(torch.bfloat16, M, N, K, a_k, b_k)


@pytest.mark.parametrize("bn", [32, 64, 128, 256])
def test_gemm_bf16_tile_widths(bn):
    _run(torch.bfloat16, 300, 520, 448, True, True, bn=bn)
    if bn >= 64:
        _run(torch.bfloat16, 300, 520, 448, False, False, bn=bn)


@pytest.mark.parametrize("M,N,K", [(700, 520, 448), (256, 256, 64), (1000, 2531, 512), (4096, 512, 2048), (130, 300, 1000)])
def test_gemm_bf16_wide_tile(M, N, K):
    """256 x 256 tile (two TMEM accumulators per CTA sharing one B stage), forced with bn = 512: ragged M / N / K, bias, tanh."""
    _run(torch.bfloat16, M, N, K, True, True, bn=512)
    _run(torch.bfloat16, M, N, K, True, True, bn=512, act=1, bias=False, seed=3)


def test_gemm_bf16_long_k_and_act():
    _run(torch.bfloat16, 512, 2048, 50176 // 8, False, False, bias=False)
    _run(torch.bfloat16, 64, 1024, 2048, True, True, act=1)
    _run(torch.float32, 64, 1024, 512, True, True, act=2)


def test_gemm_bf16_simt_crosscheck():
    from sat_b200 import _lib
    L = _lib.lib()
    L.sat_set_force_simt(1)
    try:
        _run(torch.bfloat16, 200, 72, 136, True, True)
        _run(torch.bfloat16, 200, 72, 136, False, False)
    finally:
        L.sat_set_force_simt(0)
