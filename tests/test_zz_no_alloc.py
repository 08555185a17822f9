"""GPU: `sat_attn_step_fwd_static` -- the attention step with its scratch in a `static __device__` array inside the library
(no counterpart in the reference; added for callers that cannot allocate) -- against
the existing `sat_attn_step_fwd` with a caller-owned scratch buffer, through the C ABI, on two calls in a row.

Both run the same streaming kernels on the same inputs, so the outputs must agree to the tolerances tests/test_gpu_attention.py
uses (alpha rtol 2e-4 / atol 1e-6 fp32, rtol 5e-2 / atol 2e-3 bf16; context rtol 2e-4 / atol 2e-5 fp32, 3e-2 bf16) -- in fact bit
for bit, which is asserted too.  The batch sizes are small on purpose: the work partition then CUTS images, so the partial records
in the scratch and the combine kernel are really exercised (with whole images per CTA the scratch is never read).
The second call uses other inputs and another batch size: the array keeps the first call's contents and must not care.
This file sorts last (pytest -x must not let a failure here hide other tests).
"""
import pytest
import torch

pytestmark = pytest.mark.gpu

TOL_ALPHA = {torch.float32: dict(rtol=2e-4, atol=1e-6), torch.bfloat16: dict(rtol=5e-2, atol=2e-3)}
TOL_CTX = {torch.float32: dict(rtol=2e-4, atol=2e-5), torch.bfloat16: dict(rtol=3e-2, atol=3e-2)}


def _inputs(NI, NB, L, D, H, dt, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    R = NI * NB
    a = torch.relu(torch.randn(NI, L, D, device="cuda", generator=g)).to(dt).contiguous()
    wa = (torch.randn(NI, L, H, device="cuda", generator=g) * 0.5).to(dt).contiguous()
    uf = torch.randn(R, H + D, device="cuda", generator=g).contiguous()
    v = (torch.randn(H, device="cuda", generator=g) * 0.05).contiguous()
    bv = torch.zeros(1, device="cuda")
    return a, wa, uf, v, bv


def _outputs(R, L, D, dt):
    # poisoned: whatever a call does not write shows up as NaN
    return (torch.full((R, L), float("nan"), device="cuda"), torch.full((R, D), float("nan"), device="cuda", dtype=dt),
            torch.full((R, D), float("nan"), device="cuda", dtype=dt), torch.full((R, D), float("nan"), device="cuda"))


def _call(static, dt, NI, NB, L, D, H, inp, scratch=None):
    from sat_b200 import _lib
    Lb = _lib.lib()
    a, wa, uf, v, bv = inp
    alpha, ctx, z, c32 = _outputs(NI * NB, L, D, dt)
    st = torch.cuda.current_stream().cuda_stream
    head = (_lib.dtype_code(dt), NI, NB, L, D, H, a.data_ptr(), wa.data_ptr(), uf.data_ptr(), H + D, v.data_ptr(), bv.data_ptr(), 1,
            alpha.data_ptr(), L, ctx.data_ptr(), D, z.data_ptr(), D, c32.data_ptr(), D)
    if static:
        _lib.check(Lb.sat_attn_step_fwd_static(*head, st), "sat_attn_step_fwd_static")
    else:
        _lib.check(Lb.sat_attn_step_fwd(*head, scratch.data_ptr(), scratch.numel(), st), "sat_attn_step_fwd")
    torch.cuda.synchronize()
    return alpha, ctx, z, c32


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16], ids=["fp32", "bf16"])
def test_static_scratch_variant_matches_existing_entry_point_twice(dt):
    from sat_b200 import _lib
    # (NI, NB, L, D, H): cut partitions (few images), a beam-shaped call, and a whole-image partition in between
    calls = [(5, 1, 196, 2048, 512, 21), (3, 5, 196, 512, 512, 22), (70, 1, 196, 512, 768, 23), (2, 1, 49, 2048, 768, 24)]
    for NI, NB, L, D, H, seed in calls:
        inp = _inputs(NI, NB, L, D, H, dt, seed)
        scratch = _lib.attn_scratch(NI, NB, L, D, H, "cuda")
        assert scratch is not None
        scratch.fill_(0xFF)
        ref = _call(False, dt, NI, NB, L, D, H, inp, scratch)
        free_before = torch.cuda.mem_get_info()[0]
        got = _call(True, dt, NI, NB, L, D, H, inp)
        # driver-level free memory: the call's own (small) output tensors come out of torch's cached blocks or at most one new
        # 2 MiB small-pool segment; a cudaMalloc of the scratch inside the library would show up on top of that
        assert free_before - torch.cuda.mem_get_info()[0] <= 2 * 1024 * 1024
        for name, g_, r_, tol in (("alpha", got[0], ref[0], TOL_ALPHA[dt]), ("ctx", got[1], ref[1], TOL_CTX[dt]),
                                  ("z", got[2], ref[2], TOL_CTX[dt]), ("ctx_f32", got[3], ref[3], TOL_CTX[torch.float32])):
            assert torch.isfinite(r_.float()).all() and torch.isfinite(g_.float()).all(), name
            torch.testing.assert_close(g_.float(), r_.float(), **tol, msg=lambda m: f"{name} (NI={NI}, NB={NB}, D={D}): {m}")
            assert torch.equal(g_, r_), f"{name}: same kernels, same inputs -> identical bits expected"
        assert (got[0].sum(1) - 1).abs().max().item() < 1e-5


def test_static_scratch_variant_rejects_what_does_not_fit():
    from sat_b200 import _lib
    Lb = _lib.lib()
    dt = torch.bfloat16
    # beyond the maximum: 1000 images at D = 2048 need 2.4 MB of records + 8 MB of per-image rows > 4 MiB
    NI, NB, L, D, H = 1000, 1, 4, 2048, 512
    inp = _inputs(NI, NB, L, D, H, dt, 3)
    alpha, ctx, z, c32 = _outputs(NI, L, D, dt)
    a, wa, uf, v, bv = inp
    rc = Lb.sat_attn_step_fwd_static(_lib.dtype_code(dt), NI, NB, L, D, H, a.data_ptr(), wa.data_ptr(), uf.data_ptr(), H + D,
                                     v.data_ptr(), bv.data_ptr(), 1, alpha.data_ptr(), L, ctx.data_ptr(), D, z.data_ptr(), D,
                                     c32.data_ptr(), D, torch.cuda.current_stream().cuda_stream)
    msg = Lb.sat_last_error().decode()
    assert rc != 0 and "SAT_STATIC_SCRATCH_BYTES" in msg and "needs" in msg, msg
    torch.cuda.synchronize()
    assert torch.isnan(alpha).all()                       # nothing was launched
    # a shape the streaming kernels do not serve (it needs no scratch at all) is refused with a pointer to the other entry point
    rc = Lb.sat_attn_step_fwd_static(_lib.dtype_code(dt), 2, 1, 4, 64, 64, a.data_ptr(), wa.data_ptr(), uf.data_ptr(), 128,
                                     v.data_ptr(), bv.data_ptr(), 1, alpha.data_ptr(), 4, ctx.data_ptr(), 64, z.data_ptr(), 64,
                                     None, 0, torch.cuda.current_stream().cuda_stream)
    assert rc != 0 and "sat_attn_step_fwd" in Lb.sat_last_error().decode()
