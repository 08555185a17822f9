"""GPU: the sizes at the edges of what the reference accepts, through the module surface (C ABI underneath) against the oracle.

The reference has no test suite; these are the degenerate shapes its own code paths admit: one caption, one decode step,
captions that hold only <start> (decoder.py:92-96: the loop never runs), one annotation vector (softmax over a single location),
ragged sizes that are no multiple of any tile, a vocabulary smaller than one warp, beam 1, and a beam wider than the
vocabulary (decoder.py:208 `topk(beam_size)` raises).  fp32 tolerances are those of tests/test_gpu_decoder.py.
"""
import pytest
import torch

from oracle import sat_oracle as O
from sat_b200 import synth

pytestmark = pytest.mark.gpu

TP, TA = dict(rtol=2e-4, atol=1e-4), dict(rtol=2e-4, atol=1e-6)


def _decoder(V, D, P, **kw):
    from sat_b200.decoder import Decoder
    dec = Decoder(V, D, tf=kw.get("tf", True), ado=kw.get("ado", False), attention=kw.get("attention", True),
                  compute_dtype=kw.get("dt", torch.float32), gate=kw.get("gate", True))
    dec.load_state_dict(P, strict=True)
    return dec.cuda().eval()


EDGE = [
    # B, L, D, T, V, tf, ado
    (1, 196, 512, 1, 2531, True, False),      # one caption, one step
    (1, 1, 512, 3, 50, True, False),          # one annotation vector: alpha == 1
    (3, 7, 24, 2, 5, True, True),             # nothing is a multiple of a tile; vocabulary of 5
    (2, 49, 2048, 1, 31, False, False),       # free-running with a single step: the <start> embedding only
    (7, 13, 520, 5, 33, False, True),         # ragged, free-running, advanced deep output
    (149, 4, 512, 2, 40, True, False),        # one more image than SMs with a 4-row item
]


@pytest.mark.parametrize("B,L,D,T,V,tf,ado", EDGE, ids=lambda v: str(int(v)))
def test_degenerate_shapes_forward_backward_vs_oracle(B, L, D, T, V, tf, ado):
    P = synth.make_params(V, D, 512, ado=ado, seed=41, emb_scale=0.5)
    a, cap = synth.make_inputs(B, L, D, T, V, seed=41)
    dec = _decoder(V, D, P, tf=tf, ado=ado)
    Pr = {k: v.clone().requires_grad_(True) for k, v in P.items()}
    preds, alphas = dec(a.cuda(), cap.cuda())
    pr, ar = O.decoder_forward(Pr, a, cap, tf=tf, ado=ado, attention=True, gate=True)
    assert preds.shape == (B, T, V) and alphas.shape == (B, T, L)
    torch.testing.assert_close(alphas.cpu(), ar.detach(), **TA)
    torch.testing.assert_close(preds.cpu(), pr.detach(), **TP)
    assert (alphas.sum(2) - 1).abs().max().item() < 1e-5
    if L == 1:
        assert torch.equal(alphas, torch.ones_like(alphas))
    O.train_loss(preds, alphas, cap.cuda()).backward()
    O.train_loss(pr, ar, cap).backward()
    unused = dec.unused_parameter_keys()
    for n, p in dec.named_parameters():
        if n in unused:
            assert p.grad is None, n
            continue
        ref, got = Pr[n].grad, p.grad.cpu()
        scale = ref.abs().max().item() + 1e-12
        if n == "attention.v.bias" or (L == 1 and n.startswith("attention.")):
            # softmax over one location (or any shift) has zero gradient: both sides hold rounding noise only
            assert got.abs().max().item() < 1e-5, (n, got.abs().max().item())
            continue
        assert (got - ref).abs().max().item() <= 2e-3 * scale + 1e-7, (n, (got - ref).abs().max().item(), scale)


@pytest.mark.parametrize("B,width", [(4, 1), (0, 6), (0, 1)], ids=["start_only", "empty_batch", "both"])
def test_empty_sequence_and_empty_batch_like_reference(B, width):
    """decoder.py:92-96: preds / alphas are allocated as zeros of (B, max_timespan, ...) and the loop body never runs."""
    V, D, L = 60, 512, 49
    P = synth.make_params(V, D, 512, seed=5, emb_scale=0.5)
    dec = _decoder(V, D, P)
    a = torch.randn(B, L, D, device="cuda")
    cap = torch.zeros(B, width, dtype=torch.long, device="cuda")
    preds, alphas = dec(a, cap)
    T = width - 1
    assert preds.shape == (B, T, V) and alphas.shape == (B, T, L)
    assert preds.dtype == torch.float32 and preds.device.type == "cuda"
    if B:
        pr, ar = O.decoder_forward(P, a.cpu(), cap.cpu(), tf=True, ado=False, attention=True, gate=True)
        assert pr.shape == preds.shape and ar.shape == alphas.shape


def test_int32_and_strided_captions_are_accepted():
    V, D, B, L, T = 90, 512, 5, 49, 4
    P = synth.make_params(V, D, 512, seed=6, emb_scale=0.5)
    a, cap = synth.make_inputs(B, L, D, T, V, seed=6)
    dec = _decoder(V, D, P)
    base, base_al = dec(a.cuda(), cap.cuda())
    wide = torch.zeros(B, 2 * (T + 1), dtype=torch.long)
    wide[:, ::2] = cap
    for c in (cap.to(torch.int32).cuda(), wide.cuda()[:, ::2], cap):          # int32, column-strided view, host tensor
        got, got_al = dec(a.cuda(), c)
        assert torch.equal(got, base) and torch.equal(got_al, base_al)


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16], ids=["fp32", "bf16"])
def test_beam_one_and_small_vocabulary(dt):
    """beam 1 is greedy decoding (decoder.py:160-269 with k = 1); fp32 tokens equal the oracle's bit for bit."""
    V, D, L, NI = 12, 512, 49, 9
    P = synth.make_params(V, D, 512, seed=8, emb_scale=0.5)
    a, _ = synth.make_inputs(NI, L, D, 2, V, seed=8)
    dec = _decoder(V, D, P, dt=dt)
    for beam in (1, 3, V):
        res = dec.caption_batch(a.cuda(), beam, max_steps=12)
        assert len(res) == NI
        for i in range(NI):
            sent, al = res[i]
            assert len(sent) >= 1 and all(0 <= int(w) < V for w in sent)
            if dt == torch.float32:
                ref_sent, _, _ = O.caption(P, a[i], beam, ado=False, attention=True, max_steps=12)
                assert [int(w) for w in sent] == [int(w) for w in ref_sent], (beam, i)


def test_beam_wider_than_vocabulary_is_refused():
    """decoder.py:208: `top_preds.view(-1).topk(beam_size)` on the V entries of the first step raises for beam > V."""
    V, D, L = 4, 512, 49
    P = synth.make_params(V, D, 512, seed=8, emb_scale=0.5)
    dec = _decoder(V, D, P)
    a = torch.randn(2, L, D, device="cuda")
    with pytest.raises((RuntimeError, ValueError)):
        dec.caption_batch(a, V + 1, max_steps=5)
    torch.cuda.synchronize()
    dec.caption_batch(a, V, max_steps=5)                  # the library is usable after the refusal


def test_bad_shapes_raise_before_any_launch():
    V, D, L = 30, 512, 49
    P = synth.make_params(V, D, 512, seed=8, emb_scale=0.5)
    dec = _decoder(V, D, P)
    cap = torch.zeros(2, 3, dtype=torch.long, device="cuda")
    with pytest.raises(RuntimeError):
        dec(torch.randn(2, L, D + 8, device="cuda"), cap)                    # wrong encoder_dim
    with pytest.raises(RuntimeError):
        dec(torch.randn(2, L * D, device="cuda"), cap)                       # not (B, L, D)
    with pytest.raises(TypeError):
        dec(torch.randn(2, L, D, device="cuda").half(), cap)                 # fp16 features are not a supported storage type
    with pytest.raises(Exception):
        dec(torch.randn(2, L, D), cap)                                       # host features: the product path has no CPU fallback
