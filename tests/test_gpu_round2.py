"""GPU: surface gaps closed in round 2, each against the oracle (or the reference's documented behaviour):
  * gradient wrt img_features (train.py:132 leaves the encoder outside no_grad) -- generic + streaming attention kernels,
    no-attention branch, advanced deep output;
  * fused loss with the advanced deep output (the flag set every train_models.py entry uses);
  * out-of-range caption ids raise IndexError like nn.Embedding / CrossEntropyLoss;
  * sequence_accuracy on tied logits (final ReLU of ado: blocks of zeros) admits at most k ties, NaN never counts;
  * calculate_caption_lengths / greedy tokens / feature-cache gather kernels;
  * FusedTrainStep under CUDA-graph replay == eager; StepLR; parameter-storage guard.
"""
import pytest
import torch

from oracle import sat_oracle as O
from sat_b200 import synth

pytestmark = pytest.mark.gpu


def _dec(cfg, P, dt=torch.float32):
    from sat_b200.decoder import Decoder
    dec = Decoder(cfg["V"], cfg["D"], tf=cfg.get("tf", True), ado=cfg.get("ado", False), bert=False,
                  attention=cfg.get("attention", True), compute_dtype=dt, gate=cfg.get("gate", True))
    dec.load_state_dict(P, strict=True)
    return dec.cuda().eval()


DFEAT_CASES = [
    dict(V=300, D=512, B=5, L=196, T=4),                                   # streaming kernels, cut partition
    dict(V=300, D=2048, B=66, L=196, T=3),                                 # streaming kernels, whole items
    dict(V=300, D=64, B=4, L=10, T=3),                                     # generic cluster kernels (da accumulated in-kernel)
    dict(V=300, D=512, B=4, L=49, T=3, attention=False),                   # z = mean feature at every step
    dict(V=300, D=512, B=6, L=196, T=3, ado=True),                         # + un-gated context path of the deep output
    dict(V=300, D=512, B=4, L=49, T=3, attention=False, ado=True),
    dict(V=300, D=512, B=6, L=196, T=3, gate=False),
    dict(V=300, D=512, B=5, L=196, T=3, tf=False),
]


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16], ids=["fp32", "bf16"])
@pytest.mark.parametrize("case", DFEAT_CASES, ids=lambda c: "D%d_B%d_L%d_att%d_ado%d_g%d_tf%d" % (
    c["D"], c["B"], c["L"], c.get("attention", True), c.get("ado", False), c.get("gate", True), c.get("tf", True)))
def test_img_features_gradient_vs_oracle(case, dt):
    cfg = dict(case)
    P = synth.make_params(cfg["V"], cfg["D"], 512, ado=cfg.get("ado", False), seed=5, emb_scale=0.5)
    a, cap = synth.make_inputs(cfg["B"], cfg["L"], cfg["D"], cfg["T"], cfg["V"], seed=5)
    dec = _dec(cfg, P, dt)
    feats = a.cuda().requires_grad_(True)
    preds, alphas = dec(feats, cap.cuda())
    O.train_loss(preds, alphas, cap.cuda()).backward()
    Pr = {k: v.clone().requires_grad_(True) for k, v in P.items()}
    ar_in = a.clone().requires_grad_(True)
    if dt != torch.float32:
        # the library casts fp32 features to the compute dtype once (Q20 hand-off); that input quantisation is part of the
        # bf16 mode's definition, not kernel error: the oracle differentiates through the same rounded features
        ar_in = a.to(dt).float().requires_grad_(True)
    # bf16 through the ReLUs of the advanced deep output: compare with the oracle that rounds what the kernels store in bf16
    # at the same points (mask flips are quantisation, not kernel error; see tests/test_gpu_fullsize.py)
    relu_bf16 = dt != torch.float32 and cfg.get("ado", False)
    with O.emulate_storage(dt if relu_bf16 else None):
        pr, ar = O.decoder_forward(Pr, ar_in, cap, tf=cfg.get("tf", True), ado=cfg.get("ado", False),
                                   attention=cfg.get("attention", True), gate=cfg.get("gate", True),
                                   relu_masks=(preds.detach().cpu() > 0) if relu_bf16 else None)
    O.train_loss(pr, ar, cap).backward()
    got, ref = feats.grad.cpu(), ar_in.grad
    assert got.shape == ref.shape and got.dtype == torch.float32
    rel = (got - ref).norm().item() / (ref.norm().item() + 1e-30)
    if dt == torch.float32:
        assert rel <= 1e-3, rel
        assert (got - ref).abs().max().item() <= 2e-3 * ref.abs().max().item() + 1e-9
    else:
        # bf16 through free-running argmax feedback may pick another token on a near-tie: compare only when the tokens agree
        if not cfg.get("tf", True) and not torch.equal(preds.argmax(2).cpu(), pr.argmax(2)):
            pytest.skip("bf16 greedy trajectory differs from the fp32 oracle's on a near-tie")
        assert rel <= 5e-2, rel
    # parameter gradients are unchanged by asking for d img_features
    dec2 = _dec(cfg, P, dt)
    p2, a2 = dec2(a.cuda(), cap.cuda())
    O.train_loss(p2, a2, cap.cuda()).backward()
    for (k, p), (_, q) in zip(dec.named_parameters(), dec2.named_parameters()):
        if p.grad is None:
            assert q.grad is None, k
        elif k != "embedding.weight":          # scatter-add uses fp32 atomics
            assert torch.equal(p.grad, q.grad), k


def test_img_features_gradient_bf16_input_dtype():
    cfg = dict(V=200, D=512, B=3, L=196, T=3)
    P = synth.make_params(cfg["V"], cfg["D"], 512, seed=6, emb_scale=0.5)
    a, cap = synth.make_inputs(cfg["B"], cfg["L"], cfg["D"], cfg["T"], cfg["V"], seed=6)
    dec = _dec(cfg, P, torch.bfloat16)
    feats = a.cuda().to(torch.bfloat16).requires_grad_(True)
    preds, alphas = dec(feats, cap.cuda())
    O.train_loss(preds, alphas, cap.cuda()).backward()
    assert feats.grad is not None and feats.grad.dtype == torch.bfloat16 and feats.grad.shape == feats.shape
    assert torch.isfinite(feats.grad.float()).all() and feats.grad.float().abs().sum().item() > 0


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16], ids=["fp32", "bf16"])
@pytest.mark.parametrize("attention", [True, False])
def test_fused_loss_with_advanced_deep_output(dt, attention):
    """FusedTrainStep with ado=True (train_models.py passes --ado in every entry): loss and gradients vs the oracle."""
    from sat_b200.decoder import Decoder
    from sat_b200.trainer import FusedTrainStep
    V, D, B, L, T = 1003, 512, 40, 196, 5
    P = synth.make_params(V, D, 512, ado=True, seed=13, emb_scale=0.5)
    a, cap = synth.make_inputs(B, L, D, T, V, seed=13)
    masks = synth.make_dropout_masks(B, T, 512)
    dec = Decoder(V, D, tf=True, ado=True, bert=False, attention=attention, compute_dtype=dt)
    dec.load_state_dict(P, strict=True)
    dec = dec.cuda().train()
    tr = FusedTrainStep(dec, lr=1e-3, distributed=False)
    loss = tr.step(a.cuda().to(dt), cap.cuda(), dropout_mask=masks.cuda(), optimizer_step=False)
    torch.cuda.synchronize()
    tr.check_status()
    Pr = {k: v.clone().requires_grad_(True) for k, v in P.items()}
    pr, ar = O.decoder_forward(Pr, a, cap, tf=True, ado=True, attention=attention, dropout_masks=masks)
    loss_r = O.train_loss(pr, ar, cap, 1.0)
    loss_r.backward()
    rtol = 2e-5 if dt == torch.float32 else 3e-2
    assert abs(loss[2].item() - loss_r.item()) <= rtol * abs(loss_r.item())
    assert "deep_output.weight" not in tr.train_keys and "f_out.weight" in tr.train_keys        # Q7
    for k in tr.train_keys:
        if k == "attention.v.bias":
            continue
        got, ref = tr.grad_views[k].cpu(), Pr[k].grad
        rel = (got - ref).norm().item() / (ref.norm().item() + 1e-30)
        if dt == torch.float32:
            assert rel <= 2e-3, (k, rel)
        else:
            # ReLU masks of pre-activations within bf16 noise of 0 flip: direction and size, not elementwise
            cos = torch.nn.functional.cosine_similarity(got.reshape(1, -1), ref.reshape(1, -1)).item()
            assert cos > 0.985 and rel <= 0.15, (k, cos, rel)
    # the unfused path (preds through autograd + torch loss) gives the same gradients
    dec2 = Decoder(V, D, tf=True, ado=True, bert=False, attention=attention, compute_dtype=dt)
    dec2.load_state_dict(P, strict=True)
    dec2 = dec2.cuda().train()
    p2, a2 = dec2(a.cuda().to(dt), cap.cuda(), dropout_masks=masks.cuda())
    O.train_loss(p2, a2, cap.cuda()).backward()
    named2 = dict(dec2.named_parameters())
    for k in tr.train_keys:
        if k in ("attention.v.bias", "embedding.weight"):
            continue
        g1, g2 = tr.grad_views[k], named2[k].grad
        assert (g1 - g2).abs().max().item() <= (1e-4 if dt == torch.float32 else 2e-2) * (g2.abs().max().item() + 1e-12), k


def test_out_of_range_tokens_raise_index_error():
    from sat_b200.trainer import FusedTrainStep
    cfg = dict(V=100, D=512, B=3, L=49, T=4)
    P = synth.make_params(cfg["V"], cfg["D"], 512, seed=2, emb_scale=0.5)
    a, cap = synth.make_inputs(cfg["B"], cfg["L"], cfg["D"], cfg["T"], cfg["V"], seed=2)
    dec = _dec(cfg, P)
    bad = cap.clone()
    bad[1, 2] = cfg["V"] + 7
    with pytest.raises(IndexError):
        dec(a.cuda(), bad.cuda())
    neg = cap.clone()
    neg[0, 1] = -1
    with pytest.raises(IndexError):
        dec.forward_loss(a.cuda(), neg.cuda())
    # the last column is a target the loss drops and never an input (Q10): accepted, like upstream
    last = cap.clone()
    last[:, -1] = cfg["V"] + 1
    dec(a.cuda(), last.cuda())
    # the fused trainer flags it on the device and raises when asked
    tr = FusedTrainStep(dec.train(), distributed=False)
    tr.step(a.cuda(), bad.cuda())
    with pytest.raises(IndexError):
        tr.check_status()
    tr.step(a.cuda(), cap.cuda())
    tr.check_status()


def test_sequence_accuracy_ties_and_nan():
    from sat_b200 import metrics
    V = 50
    preds = torch.zeros(2, 3, V, device="cuda")                 # all-equal rows: every logit ties with the target's
    preds[0, 0, 7] = 1.0
    targets = torch.tensor([[7, 0, 9], [4, 49, 3]], device="cuda")
    rank = metrics.logit_rank(preds, targets).cpu()
    # row (0,0): target 7 is the unique maximum -> 0; all-equal rows: rank = target index (lower ids first)
    assert rank.tolist() == [[0, 0, 9], [4, 49, 3]]
    # top-5 on an all-zero row admits exactly the five lowest ids, as many as topk(5) can return
    hits = sum(int(metrics.logit_rank(torch.zeros(1, 1, V, device="cuda"), torch.tensor([[t]], device="cuda")).item() < 5) for t in range(V))
    assert hits == 5
    # ReLU'd logits (ado): target logit 0 with three positive entries elsewhere
    row = torch.zeros(1, 1, V, device="cuda")
    row[0, 0, [10, 20, 30]] = torch.tensor([3.0, 2.0, 1.0], device="cuda")
    assert metrics.logit_rank(row, torch.tensor([[0]], device="cuda")).item() == 3        # after the three positives
    assert metrics.logit_rank(row, torch.tensor([[40]], device="cuda")).item() == 3 + 37  # + the 37 zeros with a lower id
    # NaN target logit / NaN row: never correct
    nan = torch.full((1, 1, V), float("nan"), device="cuda")
    assert metrics.logit_rank(nan, torch.tensor([[5]], device="cuda")).item() == V
    assert metrics.sequence_accuracy(nan, torch.tensor([[5]], device="cuda"), 1, ignore_index=3) == 0
    # agreement with the reference formula where there are no ties
    g = torch.Generator(device="cuda").manual_seed(0)
    p = torch.randn(4, 6, V, device="cuda", generator=g)
    t = torch.randint(0, V, (4, 6), device="cuda", generator=g)
    for k in (1, 5):
        assert abs(metrics.sequence_accuracy(p, t, k, ignore_index=3) - O.sequence_accuracy(p.cpu(), t.cpu(), k, ignore_index=3)) < 1e-9


def test_caption_token_count_and_greedy_tokens():
    from sat_b200 import metrics
    _, cap = synth.make_inputs(37, 4, 8, 26, 2531, seed=3)
    skip = torch.tensor([3, 0, 1])                                           # pad, start, eos (train.py:177)
    want = int((~cap.unsqueeze(-1).eq(skip).any(-1)).sum().item())           # utils.py:101-107
    assert metrics.calculate_caption_lengths(cap.cuda(), skip) == want
    assert metrics.calculate_caption_lengths(cap.cuda()[:, ::2], [3, 0, 1]) == int((~cap[:, ::2].unsqueeze(-1).eq(skip).any(-1)).sum())
    g = torch.Generator(device="cuda").manual_seed(1)
    preds = torch.randn(9, 7, 1003, device="cuda", generator=g)
    preds[0, 0, 5] = preds[0, 0, 900] = 50.0                                 # tie: first maximum wins (torch.max)
    got = metrics.greedy_tokens(preds)
    assert torch.equal(got, torch.max(preds, dim=2)[1]) and got[0, 0].item() == 5


def test_feature_cache_gather_bit_exact():
    from sat_b200.feature_cache import FeatureCache
    N, L, D = 40, 196, 512
    g = torch.Generator().manual_seed(0)
    feats = torch.relu(torch.randn(N, L, D, generator=g))
    fc = FeatureCache(N, L, D, dtype=torch.bfloat16, device="cuda")
    fc.put(torch.arange(0, 24), feats[:24].cuda())
    # the encoder's permuted view (Q20) is accepted as is
    perm = feats[24:].cuda().permute(0, 2, 1).contiguous().permute(0, 2, 1)
    fc.put(torch.arange(24, N), perm)
    idx = torch.tensor([3, 39, 0, 3, 17, 24], device="cuda")
    out = fc.gather(idx)
    assert out.dtype == torch.bfloat16 and out.shape == (6, L, D) and out.is_contiguous()
    assert torch.equal(out.cpu(), feats[idx.cpu()].to(torch.bfloat16))
    fc.check()
    buf = torch.empty(6, L, D, device="cuda", dtype=torch.bfloat16)
    assert fc.gather(idx, out=buf).data_ptr() == buf.data_ptr() and torch.equal(buf, out)
    fc.gather(torch.tensor([1, N + 2], device="cuda"))
    with pytest.raises(IndexError):
        fc.check()


def test_cuda_graph_step_equals_eager_and_steplr():
    from sat_b200.decoder import Decoder
    from sat_b200.trainer import FusedTrainStep
    V, D, B, L, T = 500, 512, 48, 196, 5
    P = synth.make_params(V, D, 512, seed=17, emb_scale=0.5)
    a, cap = synth.make_inputs(B, L, D, T, V, seed=17)
    ad, capd = a.cuda().to(torch.bfloat16), cap.cuda()
    res = {}
    for graph in (False, True):
        dec = Decoder(V, D, tf=True, attention=True, compute_dtype=torch.bfloat16)
        dec.load_state_dict(P, strict=True)
        dec = dec.cuda().eval()                       # no dropout: eager and replayed steps see the same inputs
        tr = FusedTrainStep(dec, lr=1e-3, distributed=False, step_size=2, gamma=0.5, use_cuda_graph=graph)
        losses = []
        for it in range(7):
            losses.append(tr.step(ad, capd)[2].item())
            if it in (2, 4):
                tr.scheduler_step()
        torch.cuda.synchronize()
        res[graph] = (losses, tr.flat_param.clone(), tr.step_dev.item(), tr.lr, tr.graph_replays)
    assert res[True][4] == 5 and res[False][4] == 0            # two eager warm-up steps, then capture + replays
    assert res[True][2] == res[False][2] == 7
    assert abs(res[True][3] - 1e-3 * 0.5) < 1e-12              # two epochs, step_size 2: one decay
    # identical kernels on identical inputs; only the embedding scatter-add's fp32 atomics may reorder
    for l0, l1 in zip(res[False][0], res[True][0]):
        assert abs(l0 - l1) <= 1e-4 * abs(l0)
    assert (res[True][1] - res[False][1]).abs().max().item() <= 2e-4
    assert res[False][0][-1] < res[False][0][0]                # and it trains


def test_graph_step_draws_fresh_dropout_masks():
    from sat_b200.decoder import Decoder
    from sat_b200.trainer import FusedTrainStep
    V, D, B, L, T = 300, 512, 8, 196, 4
    P = synth.make_params(V, D, 512, seed=3, emb_scale=0.5)
    a, cap = synth.make_inputs(B, L, D, T, V, seed=3)
    dec = Decoder(V, D, tf=True, attention=True, compute_dtype=torch.float32)
    dec.load_state_dict(P)
    dec = dec.cuda().train()
    tr = FusedTrainStep(dec, lr=0.0, distributed=False, use_cuda_graph=True)       # lr 0: weights fixed, only the mask changes
    ad, capd = a.cuda(), cap.cuda()                                # fixed device buffers: the graph is tied to their addresses
    vals = [tr.step(ad, capd)[0].item() for _ in range(6)]
    assert tr.graph_replays == 4
    assert len({round(v, 6) for v in vals[2:]}) > 1, vals       # replays see different masks


def test_three_bucket_update_equals_default():
    """FusedTrainStep.buckets = 3 (side-chain gradients applied on the comm stream under the main chain of the backward tail, one
    tick at the end) gives the same parameters as the default single Adam launch."""
    from sat_b200.decoder import Decoder
    from sat_b200.trainer import FusedTrainStep
    V, D, B, L, T = 400, 512, 40, 196, 4
    P = synth.make_params(V, D, 512, seed=19, emb_scale=0.5)
    a, cap = synth.make_inputs(B, L, D, T, V, seed=19)
    ad, capd = a.cuda().to(torch.bfloat16), cap.cuda()
    out = {}
    for nb in (2, 3):
        dec = Decoder(V, D, tf=True, attention=True, compute_dtype=torch.bfloat16)
        dec.load_state_dict(P, strict=True)
        tr = FusedTrainStep(dec.cuda().eval(), lr=1e-3, distributed=False)
        tr.buckets = nb
        tr.overlap_allreduce = nb == 3                 # the bucketed schedules are opt-in (one exchange / one Adam launch by default)
        for _ in range(3):
            tr.step(ad, capd)
        torch.cuda.synchronize()
        out[nb] = (tr.flat_param.clone(), int(tr.step_dev.item()))
    assert out[2][1] == out[3][1] == 3
    assert (out[2][0] - out[3][0]).abs().max().item() <= 2e-4      # embedding scatter-add atomics may reorder; Adam itself is elementwise


def test_trainer_follows_changing_batch_shapes():
    """train.py:127-140 pads every batch to ITS longest caption and the last batch of an epoch is short: T and B change from step
    to step.  A trainer that has seen (B, T) before must give the same update as a fresh one on the same step."""
    from sat_b200.decoder import Decoder
    from sat_b200.trainer import FusedTrainStep
    V, D, L = 300, 512, 49
    P = synth.make_params(V, D, 512, seed=23, emb_scale=0.5)
    shapes = [(24, 5), (24, 3), (7, 6), (24, 5)]
    batches = [synth.make_inputs(B, L, D, T, V, seed=50 + i) for i, (B, T) in enumerate(shapes)]

    def fresh():
        dec = Decoder(V, D, tf=True, attention=True, compute_dtype=torch.float32)
        dec.load_state_dict(P, strict=True)
        return FusedTrainStep(dec.cuda().eval(), lr=1e-3, distributed=False)

    tr = fresh()
    for i, (a, cap) in enumerate(batches):
        state = (tr.flat_param.clone(), tr.exp_avg.clone(), tr.exp_avg_sq.clone())
        loss = tr.step(a.cuda(), cap.cuda()).clone()
        one = fresh()                                   # the same step on a trainer that has never seen another shape
        one.flat_param.copy_(state[0])
        one.exp_avg.copy_(state[1])
        one.exp_avg_sq.copy_(state[2])
        one.step_dev.fill_(i)
        loss1 = one.step(a.cuda(), cap.cuda())
        torch.cuda.synchronize()
        tr.check_status()
        torch.testing.assert_close(loss[:3], loss1[:3], rtol=1e-6, atol=1e-6)
        torch.testing.assert_close(tr.flat_param, one.flat_param, rtol=0, atol=1e-5)      # embedding scatter-add atomics reorder


def test_param_storage_guard():
    from sat_b200.decoder import Decoder
    from sat_b200.trainer import FusedTrainStep
    V, D = 100, 512
    P = synth.make_params(V, D, 512, seed=3, emb_scale=0.5)
    a, cap = synth.make_inputs(2, 49, D, 3, V, seed=3)
    dec = Decoder(V, D, tf=True, attention=True)
    dec.load_state_dict(P)
    dec = dec.cuda()
    tr = FusedTrainStep(dec, distributed=False)
    tr.step(a.cuda(), cap.cuda())
    dec.load_state_dict(P)                                     # in-place copy: fine
    tr.step(a.cuda(), cap.cuda())
    dec.lstm.weight_ih.data = dec.lstm.weight_ih.data.clone()  # storage moved out of the flat buffer
    with pytest.raises(RuntimeError, match="flat parameter buffer"):
        tr.step(a.cuda(), cap.cuda())
