"""GPU: fused attention step (K2) forward/backward vs the oracle (attention.py:14-21 + gate)."""
import pytest
import torch

from oracle import sat_oracle as O
from sat_b200 import synth

pytestmark = pytest.mark.gpu

TOL = {torch.float32: dict(rtol=2e-4, atol=2e-5), torch.bfloat16: dict(rtol=3e-2, atol=3e-2)}


def _setup(B, L, D, H, seed=1):
    P = synth.make_params(50, D, H, seed=seed)
    a, _ = synth.make_inputs(B, L, D, 4, 50, seed=seed)
    g = torch.Generator().manual_seed(seed)
    h = torch.tanh(torch.randn(B, H, generator=g))
    return P, a, h


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("B,L,D,H", [(3, 10, 24, 512), (5, 196, 512, 512), (4, 196, 2048, 512), (2, 49, 2048, 768), (9, 7, 64, 64)])
def test_attention_module_forward_backward(dt, B, L, D, H):
    from sat_b200.attention import Attention
    P, a, h = _setup(B, L, D, H)
    att = Attention(D, H, compute_dtype=dt).cuda()
    att.load_state_dict({k[len("attention."):]: v for k, v in P.items() if k.startswith("attention.")})
    a_g = a.cuda().requires_grad_(True)
    h_g = h.cuda().requires_grad_(True)
    ctx, alpha = att(a_g, h_g)
    Pr = {k: v.clone().requires_grad_(True) for k, v in P.items() if k.startswith("attention.")}
    a_r, h_r = a.clone().requires_grad_(True), h.clone().requires_grad_(True)
    ctx_r, alpha_r = O.attention_forward(Pr, a_r, h_r)
    torch.testing.assert_close(alpha.cpu(), alpha_r, **(dict(rtol=2e-4, atol=1e-6) if dt == torch.float32 else dict(rtol=5e-2, atol=2e-3)))
    torch.testing.assert_close(ctx.float().cpu(), ctx_r, **TOL[dt])
    assert (alpha.sum(1) - 1).abs().max().item() < 1e-5          # north_star bound, both dtypes (fp32 softmax)
    g = torch.Generator().manual_seed(3)
    wc, wa = torch.randn(B, D, generator=g), torch.randn(B, L, generator=g)
    ((ctx.float() * wc.cuda()).sum() + (alpha * wa.cuda()).sum()).backward()
    ((ctx_r * wc).sum() + (alpha_r * wa).sum()).backward()
    tol = dict(rtol=1e-3, atol=1e-4) if dt == torch.float32 else dict(rtol=8e-2, atol=8e-2)
    for name, p in att.named_parameters():
        ref = Pr["attention." + name].grad
        scale = ref.abs().max().item() + 1e-6
        err = (p.grad.cpu() - ref).abs().max().item()
        assert err <= tol["atol"] * max(1.0, scale) + tol["rtol"] * scale, (name, err, scale)
    for got, ref, name in ((a_g.grad, a_r.grad, "a"), (h_g.grad, h_r.grad, "h")):
        scale = ref.abs().max().item() + 1e-6
        err = (got.cpu() - ref).abs().max().item()
        assert err <= tol["atol"] * max(1.0, scale) + tol["rtol"] * scale, (name, err, scale)


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("NB", [1, 3, 5, 10])
def test_attn_step_shared_features_rows(dt, NB):
    """Beam rows of one image share `a`/`Wa` (Decoder.caption): row r = img*NB + nb."""
    from sat_b200 import _lib
    L_ = _lib.lib()
    NI, L, D, H = 3, 196, 512, 512
    P, a, _ = _setup(NI, L, D, H, seed=5)
    g = torch.Generator().manual_seed(11)
    h = torch.tanh(torch.randn(NI * NB, H, generator=g))
    ad = a.cuda().to(dt).contiguous()
    af = ad.float().cpu()
    Wa = (af @ P["attention.W.weight"].t() + P["attention.W.bias"])
    Wad = Wa.cuda().to(dt).contiguous()
    uh = h @ P["attention.U.weight"].t() + P["attention.U.bias"]
    fb = h @ P["f_beta.weight"].t() + P["f_beta.bias"]
    uf = torch.cat((uh, fb), 1).cuda().contiguous()
    alpha = torch.empty(NI * NB, L, device="cuda")
    ctx = torch.empty(NI * NB, D, device="cuda", dtype=dt)
    z = torch.empty(NI * NB, D, device="cuda", dtype=dt)
    v = P["attention.v.weight"].reshape(-1).cuda().contiguous()
    bv = P["attention.v.bias"].cuda()
    _lib.check(L_.sat_attn_step_fwd(_lib.dtype_code(dt), NI, NB, L, D, H, ad.data_ptr(), Wad.data_ptr(), uf.data_ptr(), H + D,
                                    v.data_ptr(), bv.data_ptr(), 1, alpha.data_ptr(), L, ctx.data_ptr(), D, z.data_ptr(), D,
                                    None, 0, None, 0, torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    Wa_used = Wad.float().cpu()
    for r in range(NI * NB):
        i = r // NB
        e = torch.tanh(Wa_used[i] + uh[r]) @ P["attention.v.weight"].reshape(-1) + P["attention.v.bias"]
        al = torch.softmax(e, 0)
        c = (af[i] * al[:, None]).sum(0)
        zz = torch.sigmoid(fb[r]) * c
        torch.testing.assert_close(alpha[r].cpu(), al, **(dict(rtol=2e-4, atol=1e-6) if dt == torch.float32 else dict(rtol=5e-2, atol=2e-3)))
        torch.testing.assert_close(ctx[r].float().cpu(), c, **TOL[dt])
        torch.testing.assert_close(z[r].float().cpu(), zz, **TOL[dt])


# ---- streaming kernels (NI * groups >= 64, D in {512, 2048}, H in {512, 768}) -------------------------------------
def _ref_step(P, a_used, wa_used, uh, fb, gate=True):
    """fp32 reference of one attention step on already-projected features (rows share image r // NB)."""
    v = P["attention.v.weight"].reshape(-1)
    e = torch.tanh(wa_used + uh[:, None, :]) @ v + P["attention.v.bias"]
    al = torch.softmax(e, 1)
    c = (a_used * al[:, :, None]).sum(1)
    return al, c, (torch.sigmoid(fb) * c if gate else c)


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16], ids=["fp32", "bf16"])
@pytest.mark.parametrize("NI,NB,L,D,H", [(70, 1, 196, 512, 512), (64, 1, 196, 2048, 512), (66, 1, 49, 2048, 768), (65, 1, 197, 512, 768),
                                         (64, 5, 196, 512, 512), (33, 3, 196, 512, 512), (16, 10, 196, 512, 512),
                                         (1, 1, 196, 2048, 512), (3, 1, 5, 512, 512), (7, 5, 196, 512, 768), (300, 1, 196, 512, 512)])
def test_streaming_fwd_vs_oracle_and_generic(dt, NI, NB, L, D, H):
    from sat_b200 import _lib
    L_ = _lib.lib()
    P, a, _ = _setup(NI, L, D, H, seed=7)
    g = torch.Generator().manual_seed(13)
    R = NI * NB
    h = torch.tanh(torch.randn(R, H, generator=g))
    ad = a.cuda().to(dt).contiguous()
    af = ad.float().cpu()
    Wad = (af @ P["attention.W.weight"].t() + P["attention.W.bias"]).cuda().to(dt).contiguous()
    uh = h @ P["attention.U.weight"].t() + P["attention.U.bias"]
    fb = h @ P["f_beta.weight"].t() + P["f_beta.bias"]
    uf = torch.cat((uh, fb), 1).cuda().contiguous()
    v = P["attention.v.weight"].reshape(-1).cuda().contiguous()
    bv = P["attention.v.bias"].cuda()
    out = {}
    scratch = _lib.attn_scratch(NI, NB, L, D, H, "cuda")
    assert scratch is not None                      # these shapes are served by the streaming kernels
    scratch.fill_(0xFF)                             # the scratch needs no initialisation (NaN bit patterns here)
    # mode 1 / 2: streaming kernels with whole-item / equal-slot-count partition (cut items + combine kernel); 0: generic
    for mode in (1, 2, 0):
        alpha = torch.zeros(R, L, device="cuda")
        ctx = torch.zeros(R, D, device="cuda", dtype=dt)
        z = torch.zeros(R, D, device="cuda", dtype=dt)
        c32 = torch.zeros(R, D, device="cuda")
        extra = (c32.data_ptr(), D, scratch.data_ptr(), scratch.numel()) if mode else (None, 0, None, 0)
        L_.sat_set_attn_partition(mode)
        try:
            for rep in range(2 if mode else 1):     # twice: a launch must not depend on what the previous one left behind
                _lib.check(L_.sat_attn_step_fwd(_lib.dtype_code(dt), NI, NB, L, D, H, ad.data_ptr(), Wad.data_ptr(), uf.data_ptr(), H + D,
                                                v.data_ptr(), bv.data_ptr(), 1, alpha.data_ptr(), L, ctx.data_ptr(), D, z.data_ptr(), D,
                                                *extra, torch.cuda.current_stream().cuda_stream))
                torch.cuda.synchronize()
        finally:
            L_.sat_set_attn_partition(0)
        if mode:
            torch.testing.assert_close(c32.cpu(), ctx.float().cpu(), **TOL[dt])
        out[mode] = (alpha.cpu(), ctx.float().cpu(), z.float().cpu())
    rows = torch.arange(R) // NB
    al, c, zz = _ref_step(P, af[rows], Wad.float().cpu()[rows], uh, fb)
    ta = dict(rtol=2e-4, atol=1e-6) if dt == torch.float32 else dict(rtol=5e-2, atol=2e-3)
    for mode in (0, 1, 2):
        alpha, ctx, z = out[mode]
        torch.testing.assert_close(alpha, al, **ta)
        torch.testing.assert_close(ctx, c, **TOL[dt])
        torch.testing.assert_close(z, zz, **TOL[dt])
        assert (alpha.sum(1) - 1).abs().max().item() < 1e-5
    # the kernels agree with each other far inside the oracle tolerance
    torch.testing.assert_close(out[0][0], out[1][0], rtol=1e-5, atol=1e-7)
    torch.testing.assert_close(out[2][0], out[1][0], rtol=1e-5, atol=1e-7)


@pytest.mark.parametrize("part", [1, 2], ids=["whole", "cut"])
@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16], ids=["fp32", "bf16"])
@pytest.mark.parametrize("B,L,D,H,gate", [(70, 196, 512, 512, 1), (64, 196, 2048, 512, 1), (66, 49, 2048, 768, 0), (65, 197, 512, 768, 1),
                                          (1, 196, 2048, 512, 1), (3, 5, 512, 512, 1), (200, 196, 512, 512, 1)])
def test_streaming_bwd_vs_autograd(dt, B, L, D, H, gate, part):
    """sat_attn_step_bwd with dpre_out (streaming kernel) vs torch autograd of the same step in fp32."""
    from sat_b200 import _lib
    L_ = _lib.lib()
    P, a, _ = _setup(B, L, D, H, seed=8)
    g = torch.Generator().manual_seed(17)
    h = torch.tanh(torch.randn(B, H, generator=g))
    ad = a.cuda().to(dt).contiguous()
    af = ad.float().cpu()
    wa_f = (af @ P["attention.W.weight"].t() + P["attention.W.bias"]).cuda().to(dt).contiguous()
    uh = (h @ P["attention.U.weight"].t() + P["attention.U.bias"])
    fb = (h @ P["f_beta.weight"].t() + P["f_beta.bias"])
    dz = torch.randn(B, D, generator=g)
    dce = torch.randn(B, D, generator=g) * 0.5
    ga = torch.randn(B, L, generator=g) * 0.1
    # autograd reference on the rounded operands
    wa_r = wa_f.float().cpu().requires_grad_(True)
    uh_r, fb_r = uh.clone().requires_grad_(True), fb.clone().requires_grad_(True)
    v_r = P["attention.v.weight"].reshape(-1).clone().requires_grad_(True)
    bv_r = P["attention.v.bias"].clone().requires_grad_(True)
    e = torch.tanh(wa_r + uh_r[:, None, :]) @ v_r + bv_r
    al = torch.softmax(e, 1)
    c = (af * al[:, :, None]).sum(1)
    zz = torch.sigmoid(fb_r) * c if gate else c
    ((zz * dz).sum() + (c * dce).sum() + (al * ga).sum()).backward()
    # device
    uf = torch.cat((uh, fb), 1).cuda().contiguous()
    alpha_d = al.detach().cuda().contiguous()
    ctx_d = c.detach().cuda().to(dt).contiguous()
    c32 = c.detach().cuda().contiguous()
    scratch = _lib.attn_scratch(B, 1, L, D, H, "cuda")
    assert scratch is not None
    scratch.fill_(0xFF)                             # no initialisation needed
    duf = torch.zeros(B, H + D, device="cuda", dtype=dt)
    dpre = torch.zeros(B, L, H, device="cuda", dtype=dt)
    dv = torch.zeros(B, H, device="cuda")
    dbv = torch.zeros(B, device="cuda")
    vd = P["attention.v.weight"].reshape(-1).cuda().contiguous()
    dzd, dced, gad = dz.cuda(), dce.cuda(), ga.cuda()
    L_.sat_set_attn_partition(part)
    try:
        _lib.check(L_.sat_attn_step_bwd(_lib.dtype_code(dt), B, L, D, H, ad.data_ptr(), wa_f.data_ptr(), uf.data_ptr(), H + D, vd.data_ptr(),
                                        alpha_d.data_ptr(), L, ctx_d.data_ptr(), D, dzd.data_ptr(), D, dced.data_ptr(), D, gad.data_ptr(), L,
                                        gate, duf.data_ptr(), H + D, None, dv.data_ptr(), dbv.data_ptr(), None, 1, dpre.data_ptr(),
                                        c32.data_ptr(), D, scratch.data_ptr(), scratch.numel(), torch.cuda.current_stream().cuda_stream))
        torch.cuda.synchronize()
    finally:
        L_.sat_set_attn_partition(0)

    def close(got, ref, name, rel):
        scale = ref.abs().max().item() + 1e-12
        err = (got - ref).abs().max().item()
        assert err <= rel * scale, (name, err, scale)

    rel = 2e-4 if dt == torch.float32 else 2e-2
    close(dpre.float().cpu(), wa_r.grad, "dpre", rel)
    close(duf[:, :H].float().cpu(), uh_r.grad, "dUh", rel)
    if gate:
        close(duf[:, H:].float().cpu(), fb_r.grad, "dfb", rel)
    else:
        assert duf[:, H:].float().abs().max().item() == 0.0
    close(dv.sum(0).cpu(), v_r.grad, "dv", rel)
    assert abs(dbv.sum().item() - bv_r.grad.item()) < 1e-4 * max(1.0, wa_r.grad.abs().sum().item() ** 0.5)
