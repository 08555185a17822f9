"""GPU: the fused train step (`FusedTrainStep`, the path bench.py times: forward + train.py:150-162 loss + backward into
the flat gradient buffer + fused Adam, logits kept in the workspace) vs the oracle + torch.optim.Adam on the same
weights, inputs and dropout masks.  fp32: loss rtol 2e-5, gradients 2e-3 of the tensor's max, parameters after two Adam
steps rtol 1e-3 of the update size; bf16: loss 3e-2, gradient direction (relative L2 error <= 5e-2)."""
import pytest
import torch

from oracle import sat_oracle as O
from sat_b200 import synth

pytestmark = pytest.mark.gpu

CASES = [
    dict(V=2531, D=512, B=9, L=196, T=5, ado=False, attention=True),      # small batch: cut partition + combine kernels
    dict(V=1003, D=2048, B=66, L=196, T=4, ado=False, attention=True),    # whole-item partition, V not a multiple of 8
    dict(V=640, D=512, B=6, L=49, T=4, ado=False, attention=False),
    dict(V=300, D=512, B=144, L=196, T=3, ado=False, attention=True),     # B >= 128: recurrent loop as two half-batches on two streams
]


def _ids(c):
    return "V%d_D%d_B%d_att%d" % (c["V"], c["D"], c["B"], c["attention"])


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16], ids=["fp32", "bf16"])
@pytest.mark.parametrize("cfg", CASES, ids=_ids)
def test_fused_train_step_vs_oracle_adam(cfg, dt):
    from sat_b200.decoder import Decoder
    from sat_b200.trainer import FusedTrainStep
    Hd = 512
    P = synth.make_params(cfg["V"], cfg["D"], Hd, ado=cfg["ado"], seed=11, emb_scale=0.5)
    a, cap = synth.make_inputs(cfg["B"], cfg["L"], cfg["D"], cfg["T"], cfg["V"], seed=11)
    masks = synth.make_dropout_masks(cfg["B"], cfg["T"], Hd)
    dec = Decoder(cfg["V"], cfg["D"], tf=True, ado=cfg["ado"], bert=False, attention=cfg["attention"], compute_dtype=dt)
    dec.load_state_dict(P, strict=True)
    dec = dec.cuda().train()
    lr = 1e-3
    tr = FusedTrainStep(dec, lr=lr, alpha_c=1.0, distributed=False)

    Pr = {k: v.clone().requires_grad_(True) for k, v in P.items()}
    opt = torch.optim.Adam(list(Pr.values()), lr=lr)
    ad, capd, md = a.cuda().to(dt), cap.cuda(), masks.cuda()
    for it in range(2):
        opt.zero_grad(set_to_none=True)
        pr, ar = O.decoder_forward(Pr, a, cap, tf=True, ado=cfg["ado"], attention=cfg["attention"], dropout_masks=masks)
        loss_r = O.train_loss(pr, ar, cap, 1.0)
        loss_r.backward()
        loss = tr.step(ad, capd, dropout_mask=md)
        torch.cuda.synchronize()
        rtol = 2e-5 if dt == torch.float32 else 3e-2
        assert abs(loss[2].item() - loss_r.item()) <= rtol * abs(loss_r.item()), (it, loss.tolist(), loss_r.item())
        if it == 0:
            for k in tr.train_keys:
                ref = Pr[k].grad
                got = tr.grad_views[k].cpu()
                if k == "attention.v.bias":
                    continue                      # true gradient is 0 (softmax shift invariance): rounding noise on both sides
                scale = ref.abs().max().item() + 1e-12
                if dt == torch.float32:
                    assert (got - ref).abs().max().item() <= 2e-3 * scale + 1e-7, k
                else:
                    rel = (got - ref).norm().item() / (ref.norm().item() + 1e-20)
                    assert rel <= 5e-2, (k, rel)
        opt.step()
    unused = dec.unused_parameter_keys()
    named = dict(dec.named_parameters())
    for k, ref in Pr.items():
        got = named[k].detach().cpu()
        if k in unused or k == "attention.v.bias":
            continue
        if dt == torch.float32:
            # Adam moves every element by ~lr per step: compare against the size of the update, not of the weight
            # (elements whose gradient is ~0 have an ill-conditioned update direction m / sqrt(v): leave them out)
            big = Pr[k].grad.abs() > 1e-3 * Pr[k].grad.abs().max()
            err = ((got - ref.detach()).abs() * big).max().item()
            assert err <= 0.05 * lr, (k, err)
        else:
            assert (got - ref.detach()).abs().max().item() <= 2.5 * lr * 2, k        # both moved at most ~lr per step


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16], ids=["fp32", "bf16"])
@pytest.mark.parametrize("B,D", [(128, 2048), (272, 512)])
def test_half_batch_pipelining_matches_single_stream(dt, B, D):
    """sat_set_decoder_split: the two-stream half-batch loop and the single-stream loop compute the same step (only the
    GEMM tile shapes differ with the row count): loss, alphas and every gradient agree to rounding (fp32 1e-5 of the
    tensor's max, bf16 2e-2), and repeated split runs reproduce themselves (no race between the streams)."""
    from sat_b200 import _lib
    from sat_b200.decoder import Decoder
    from sat_b200.trainer import FusedTrainStep
    V, Hd, L, T = 333, 512, 196, 4
    P = synth.make_params(V, D, Hd, seed=5, emb_scale=0.5)
    a, cap = synth.make_inputs(B, L, D, T, V, seed=5)
    masks = synth.make_dropout_masks(B, T, Hd).cuda()
    ad, capd = a.cuda().to(dt), cap.cuda()
    out = {}
    try:
        for mode in (0, 1, 1):
            _lib.lib().sat_set_decoder_split(mode)
            dec = Decoder(V, D, tf=True, attention=True, compute_dtype=dt)
            dec.load_state_dict(P, strict=True)
            dec = dec.cuda().train()
            tr = FusedTrainStep(dec, lr=1e-3, distributed=False)
            loss = tr.step(ad, capd, dropout_mask=masks, optimizer_step=False)
            torch.cuda.synchronize()
            res = (loss.clone(), tr.alphas.clone(), tr.flat_grad.clone())
            if mode == 1 and 1 in out:
                assert torch.equal(out[1][1], res[1])        # alphas: bit-identical
                # gradients: identical up to the fp32 atomics of the embedding scatter-add
                assert (out[1][2] - res[2]).abs().max().item() <= 1e-6 * out[1][2].abs().max().item()
            out[mode] = res
    finally:
        _lib.lib().sat_set_decoder_split(0)
    tol = 1e-5 if dt == torch.float32 else 2e-2
    for x, y in zip(out[0], out[1]):
        assert (x - y).abs().max().item() <= tol * max(1e-6, x.abs().max().item())
    assert (out[1][1].sum(2) - 1).abs().max().item() < 1e-5
