"""GPU: sequence accuracy at k (train.py:144-145, utils.py:44-80) through sat_logit_rank -- one pass over the logits --
against the reference function's golden values, the oracle on larger random cases, and from the fused train step's
workspace logits.  Integer counts: exact."""
import os

import numpy as np
import pytest
import torch

from oracle import sat_oracle as O

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "acc_sequence_accuracy.npz")


def test_sequence_accuracy_matches_reference_golden():
    from sat_b200.metrics import sequence_accuracy
    z = np.load(GOLDEN)
    for i in range(3):
        pad = int(z[f"case{i}_cfg"][3])
        preds = torch.from_numpy(z[f"case{i}_preds"]).cuda()
        targets = torch.from_numpy(z[f"case{i}_targets"]).cuda()
        got = [sequence_accuracy(preds, targets, k, ignore_index=pad) for k in (1, 3, 5)]
        np.testing.assert_allclose(got, z[f"case{i}_acc"], rtol=0, atol=1e-9)
    assert sequence_accuracy(torch.randn(2, 3, 9).cuda(), torch.zeros(2, 3, dtype=torch.long).cuda(), 1) == 0


@pytest.mark.parametrize("B,T,V,pad", [(32, 20, 2531, 3), (5, 7, 30522, 0), (16, 50, 10000, 3), (1, 1, 7, 0)])
def test_logit_rank_vs_oracle(B, T, V, pad):
    from sat_b200.metrics import accuracy_counts, logit_rank, sequence_accuracy
    g = torch.Generator().manual_seed(B * 1000 + V)
    preds = torch.randn(B, T, V, generator=g)
    targets = torch.randint(0, V, (B, T), generator=g)
    targets[:, T // 2:] = torch.where(torch.rand(B, T - T // 2, generator=g) < 0.3, torch.tensor(pad), targets[:, T // 2:])
    preds.scatter_add_(2, targets.unsqueeze(-1), 2.5 * (torch.rand(B, T, 1, generator=g) < 0.5).float())
    pc, tc = preds.cuda(), targets.cuda()
    rank = logit_rank(pc, tc)
    want = (preds > preds.gather(2, targets.unsqueeze(-1))).sum(2).to(torch.int32)
    assert torch.equal(rank.cpu(), want)
    for k in (1, 5):
        assert sequence_accuracy(pc, tc, k, ignore_index=pad) == pytest.approx(O.sequence_accuracy(preds, targets, k, pad), abs=1e-9)
    c = accuracy_counts(rank, tc, (1, 5), pad).cpu()
    assert c[2].item() == targets.ne(pad).sum().item() and c[0] <= c[1] <= c[2]


def test_out_of_range_target_is_never_correct():
    from sat_b200.metrics import logit_rank
    preds = torch.randn(3, 11).cuda()
    rank = logit_rank(preds, torch.tensor([0, 11, -1]).cuda())
    assert rank[1].item() == 11 and rank[2].item() == 11 and 0 <= rank[0].item() < 11


def test_fused_train_step_accuracy_counts_vs_materialised_preds():
    """FusedTrainStep.accuracy_counts reads the padded logits in the workspace; the same weights through Decoder.forward
    give a (B, T, V) preds tensor: both must count the same tokens (fp32 mode, eval dropout = identical logits)."""
    from sat_b200 import synth
    from sat_b200.decoder import Decoder
    from sat_b200.metrics import accuracy_counts, logit_rank
    from sat_b200.trainer import FusedTrainStep
    V, D, B, L, T = 1003, 512, 6, 49, 5          # V not a multiple of 8: padded pitch in the workspace
    P = synth.make_params(V, D, 512, seed=3, emb_scale=0.5)
    a, cap = synth.make_inputs(B, L, D, T, V, seed=3)
    dec = Decoder(V, D, tf=True, attention=True, compute_dtype=torch.float32)
    dec.load_state_dict(P)
    dec = dec.cuda().eval()
    with torch.no_grad():
        preds, _ = dec(a.cuda(), cap.cuda())
    want = accuracy_counts(logit_rank(preds, cap[:, 1:].cuda()), cap[:, 1:].cuda(), (1, 5), 3)
    tr = FusedTrainStep(dec, lr=0.0, distributed=False)
    tr.step(a.cuda(), cap.cuda(), optimizer_step=False)
    got = tr.accuracy_counts(cap.cuda(), (1, 5), ignore_index=3)
    assert torch.equal(got.cpu(), want.cpu())
    assert got[2].item() == cap[:, 1:].ne(3).sum().item()
