"""CPU: the oracle restatement (oracle/sat_oracle.py) vs the committed live-reference vectors."""
import os

import numpy as np
import pytest
import torch

from oracle import sat_oracle as O
from tests import helpers as H


@pytest.mark.parametrize("path", H.golden_files("fwd_"), ids=os.path.basename)
def test_forward_and_grads_match_reference(path):
    cfg, P, a, cap, masks, z = H.load_fwd_case(path)
    frozen = {"embedding.weight"} if cfg["bert"] else set()          # decoder.py:32-33 (Q23)
    P = {k: v.clone().requires_grad_(k not in frozen) for k, v in P.items()}
    preds, alphas = O.decoder_forward(P, a, cap, tf=cfg["tf"], ado=cfg["ado"], attention=cfg["attention"],
                                      gate=cfg["gate"], start_token=101 if cfg["bert"] else 0,
                                      dropout_masks=masks)
    step = z["preds"].shape[2]
    stride = int(np.ceil(preds.shape[2] / step)) if step != preds.shape[2] else 1
    np.testing.assert_array_equal(preds.detach().argmax(2).numpy(), z["preds_argmax"])
    np.testing.assert_allclose(preds.detach().numpy()[:, :, ::stride], z["preds"], rtol=1e-4, atol=2e-5)
    np.testing.assert_allclose(alphas.detach().numpy(), z["alphas"], rtol=1e-4, atol=1e-6)
    assert np.abs(alphas.detach().sum(2).numpy() - 1).max() < 1e-5
    if "loss" in z:
        loss = O.train_loss(preds, alphas, cap)
        np.testing.assert_allclose(loss.item(), z["loss"][0], rtol=1e-5)
        loss.backward()
        for n in z["grad_names"]:
            n = str(n)
            if n.endswith(":none"):
                g = P[n[:-5]].grad
                assert g is None or float(g.abs().sum()) == 0.0, n
                continue
            H.check_grad_summary(H.grad_summary(P[n].grad), z["grad." + n], n)


@pytest.mark.parametrize("path", H.golden_files("cap_"), ids=os.path.basename)
def test_caption_matches_reference(path):
    cfg, P, a, z = H.load_cap_case(path)
    with torch.no_grad():
        for i in range(cfg["nimg"]):
            sent, alpha, fb = O.caption(P, a[i], cfg["beam"], ado=cfg["ado"], attention=cfg["attention"],
                                        bert=cfg["bert"])
            assert sent == z[f"sent{i}"].tolist()
            assert fb == bool(z[f"fallback{i}"][0])
            got = alpha.numpy() if torch.is_tensor(alpha) else np.array(alpha, dtype=np.float32)
            np.testing.assert_allclose(got, z[f"alpha{i}"], rtol=1e-4, atol=1e-6)


def test_oracle_sequence_accuracy_matches_reference_golden():
    """oracle.sequence_accuracy (utils.py:44-80 restated) against values the live reference function returned
    (tests/golden/make_golden_accuracy.py): exact to float rounding."""
    import os
    import numpy as np
    import torch
    from oracle import sat_oracle as O
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "acc_sequence_accuracy.npz"))
    for i in range(3):
        pad = int(z[f"case{i}_cfg"][3])
        preds, targets = torch.from_numpy(z[f"case{i}_preds"]), torch.from_numpy(z[f"case{i}_targets"])
        got = [O.sequence_accuracy(preds, targets, k, ignore_index=pad) for k in (1, 3, 5)]
        np.testing.assert_allclose(got, z[f"case{i}_acc"], rtol=0, atol=1e-9)
    assert O.sequence_accuracy(torch.randn(2, 3, 9), torch.zeros(2, 3, dtype=torch.long), 1) == z["empty_acc"][0] == 0
