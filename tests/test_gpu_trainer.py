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
