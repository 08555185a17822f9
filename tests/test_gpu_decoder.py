"""GPU: Decoder.forward / backward through the C-ABI vs (a) the committed live-reference golden vectors and
(b) the oracle on seeded inputs at config-like sizes.  Tolerances: fp32 rtol 1e-4 / atol 2e-5 on logits,
alphas row-sum within 1e-5; bf16 rtol/atol 3e-2 vs the fp32 oracle."""
import os

import numpy as np
import pytest
import torch

from oracle import sat_oracle as O
from sat_b200 import synth
from tests import helpers as H

pytestmark = pytest.mark.gpu


def build_decoder(cfg, P, dt=torch.float32):
    from sat_b200.decoder import Decoder
    if cfg.get("bert"):
        H.stub_bert()
    dec = Decoder(cfg["V"], cfg["D"], tf=cfg["tf"], ado=cfg["ado"], bert=cfg.get("bert", False), attention=cfg["attention"],
                  compute_dtype=dt, gate=cfg.get("gate", True))
    dec.load_state_dict(P, strict=True)
    return dec.cuda().eval()


@pytest.mark.parametrize("path", H.golden_files("fwd_"), ids=os.path.basename)
def test_decoder_fp32_matches_reference_golden(path):
    cfg, P, a, cap, masks, z = H.load_fwd_case(path)
    dec = build_decoder(cfg, P)
    preds, alphas = dec(a.cuda(), cap.cuda(), dropout_masks=masks)
    pn, an = preds.detach().cpu().numpy(), alphas.detach().cpu().numpy()
    step = z["preds"].shape[2]
    stride = int(np.ceil(pn.shape[2] / step)) if step != pn.shape[2] else 1
    np.testing.assert_allclose(an, z["alphas"], rtol=1e-4, atol=1e-6)
    np.testing.assert_allclose(pn[:, :, ::stride], z["preds"], rtol=1e-4, atol=5e-5)
    np.testing.assert_array_equal(pn.argmax(2), z["preds_argmax"])
    assert np.abs(an.sum(2) - 1).max() < 1e-5
    if "loss" not in z:
        return
    loss = O.train_loss(preds, alphas, cap.cuda())
    np.testing.assert_allclose(loss.item(), z["loss"][0], rtol=2e-5)
    loss.backward()
    named = dict(dec.named_parameters())
    for n in z["grad_names"]:
        n = str(n)
        if n.endswith(":none"):
            assert named[n[:-5]].grad is None, n          # reference's grad=None pattern (Q6, Q7, Q23)
            continue
        H.check_grad_summary(H.grad_summary(named[n].grad), z["grad." + n], n, rtol=5e-4, atol=5e-6)


CASES = [
    dict(V=2531, D=512, B=8, L=196, T=6, tf=True, ado=False, attention=True),
    dict(V=1000, D=2048, B=6, L=196, T=5, tf=True, ado=True, attention=True),
    dict(V=777, D=512, B=5, L=49, T=4, tf=False, ado=False, attention=True),
    dict(V=500, D=512, B=4, L=196, T=4, tf=True, ado=False, attention=False),
    dict(V=640, D=512, B=33, L=196, T=5, tf=True, ado=False, attention=True, gate=False),
    # small batches above run the streaming attention kernels with the equal-slot partition + combine kernels;
    # from ~SMs/4 items the partition is whole items per CTA
    dict(V=640, D=512, B=66, L=196, T=4, tf=True, ado=False, attention=True),
    dict(V=520, D=2048, B=64, L=196, T=3, tf=True, ado=True, attention=True),
    dict(V=520, D=512, B=65, L=49, T=3, tf=True, ado=False, attention=True, gate=False),
]


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16], ids=["fp32", "bf16"])
@pytest.mark.parametrize("case", CASES, ids=lambda c: "V%d_D%d_B%d_tf%d_ado%d_att%d" % (c["V"], c["D"], c["B"], c["tf"], c["ado"], c["attention"]))
def test_decoder_vs_oracle(case, dt):
    cfg = dict(case)
    cfg.setdefault("gate", True)
    Hd = 512
    P = synth.make_params(cfg["V"], cfg["D"], Hd, ado=cfg["ado"], seed=3, emb_scale=0.5)
    a, cap = synth.make_inputs(cfg["B"], cfg["L"], cfg["D"], cfg["T"], cfg["V"], seed=3)
    masks = synth.make_dropout_masks(cfg["B"], cfg["T"], Hd)
    dec = build_decoder(cfg, P, dt)
    Pr = {k: v.clone().requires_grad_(True) for k, v in P.items()}
    feed_cap = cap
    preds, alphas = dec(a.cuda(), cap.cuda(), dropout_masks=masks)
    pr, ar = O.decoder_forward(Pr, a, feed_cap, tf=cfg["tf"], ado=cfg["ado"], attention=cfg["attention"], gate=cfg["gate"],
                               dropout_masks=masks)
    if dt == torch.float32:
        tp, ta = dict(rtol=2e-4, atol=1e-4), dict(rtol=2e-4, atol=1e-6)
    else:
        tp, ta = dict(rtol=3e-2, atol=3e-2), dict(rtol=5e-2, atol=2e-3)
    assert (alphas.sum(2) - 1).abs().max().item() < 1e-5
    if not cfg["tf"] and dt != torch.float32:
        # free-running bf16: a near-tie argmax may legitimately diverge from the fp32 oracle, after which the two
        # trajectories are incomparable.  Teacher-style check instead of a prefix comparison: the oracle is fed OUR greedy tokens,
        # then EVERY step must agree, and our token must be an argmax of the oracle's logits up to the logit tolerance.
        ours = preds.argmax(2).cpu()
        feed = torch.cat((torch.zeros(cfg["B"], 1, dtype=torch.long), ours[:, :-1]), 1)
        pr2, ar2 = O.decoder_forward(P, a, feed_cap, tf=False, ado=cfg["ado"], attention=cfg["attention"], gate=cfg["gate"],
                                     dropout_masks=masks, feed_tokens=feed)
        torch.testing.assert_close(preds.cpu(), pr2, **tp)
        torch.testing.assert_close(alphas.cpu(), ar2, **ta)
        gap = (pr2.max(2).values - pr2.gather(2, ours.unsqueeze(2)).squeeze(2))[:, :-1].max().item()
        assert gap <= 2 * tp["atol"], gap
        return
    torch.testing.assert_close(alphas.cpu(), ar.detach(), **ta)
    torch.testing.assert_close(preds.cpu(), pr.detach(), **tp)
    loss = O.train_loss(preds, alphas, cap.cuda())
    loss_r = O.train_loss(pr, ar, cap)
    loss.backward()
    loss_r.backward()
    unused = dec.unused_parameter_keys()
    for n, p in dec.named_parameters():
        ref = Pr[n].grad
        if n in unused:
            assert p.grad is None, n
            continue
        got = p.grad.cpu()
        scale = ref.abs().max().item() + 1e-12
        err = (got - ref).abs().max().item()
        if n == "attention.v.bias":
            # softmax is shift invariant: the true gradient is 0 and both sides hold only rounding noise
            assert ref.abs().max().item() < 1e-6 and got.abs().max().item() < 1e-5, (n, got, ref)
        elif dt == torch.float32:
            assert err <= 2e-3 * scale + 1e-7, (n, err, scale)
        elif cfg["ado"]:
            # bf16 vs the fp32 oracle through ReLUs: a pre-activation within bf16 noise of 0 flips its mask and changes a
            # whole gradient row by O(one term); element-wise bounds are meaningless, direction and size are not.
            cos = torch.nn.functional.cosine_similarity(got.reshape(1, -1), ref.reshape(1, -1)).item()
            ratio = got.norm().item() / (ref.norm().item() + 1e-20)
            assert cos > 0.97 and 0.9 < ratio < 1.1, (n, cos, ratio)
        else:
            rel = (got - ref).norm().item() / (ref.norm().item() + 1e-20)
            assert rel <= 5e-2 and err <= 0.1 * scale + 1e-7, (n, rel, err, scale)


def test_strided_img_features_and_bf16_input():
    """Q20: the encoder hands over a permuted (B,C,HW)->(B,L,D) view."""
    cfg = dict(V=300, D=512, B=3, L=196, T=3, tf=True, ado=False, attention=True)
    P = synth.make_params(cfg["V"], cfg["D"], 512, seed=9, emb_scale=0.5)
    a, cap = synth.make_inputs(cfg["B"], cfg["L"], cfg["D"], cfg["T"], cfg["V"], seed=9)
    dec = build_decoder(cfg, P)
    base, _ = dec(a.cuda(), cap.cuda())
    a_perm = a.cuda().permute(0, 2, 1).contiguous().permute(0, 2, 1)      # strides (L*D, 1, L)
    assert a_perm.stride() == (cfg["L"] * cfg["D"], 1, cfg["L"])
    got, _ = dec(a_perm, cap.cuda())
    torch.testing.assert_close(got, base, rtol=0, atol=0)


def test_fused_loss_matches_unfused():
    cfg = dict(V=900, D=512, B=7, L=196, T=6, tf=True, ado=False, attention=True)
    P = synth.make_params(cfg["V"], cfg["D"], 512, seed=4, emb_scale=0.5)
    a, cap = synth.make_inputs(cfg["B"], cfg["L"], cfg["D"], cfg["T"], cfg["V"], seed=4)
    dec = build_decoder(cfg, P)
    preds, alphas = dec(a.cuda(), cap.cuda())
    O.train_loss(preds, alphas, cap.cuda()).backward()
    ref = {n: p.grad.clone() for n, p in dec.named_parameters() if p.grad is not None}
    ref_loss = O.train_loss(preds, alphas, cap.cuda()).item()
    dec.zero_grad(set_to_none=True)
    loss, _, _ = dec.forward_loss(a.cuda(), cap.cuda(), alpha_c=1.0)
    assert abs(loss.item() - ref_loss) < 2e-5 * max(1.0, abs(ref_loss))
    loss.backward()
    for n, p in dec.named_parameters():
        if n in ref:
            scale = ref[n].abs().max().item() + 1e-12
            assert (p.grad - ref[n]).abs().max().item() <= 1e-3 * scale + 1e-8, n
