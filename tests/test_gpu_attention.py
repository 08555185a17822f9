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
                                    torch.cuda.current_stream().cuda_stream))
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
