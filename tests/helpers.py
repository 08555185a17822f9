"""Shared test helpers: golden loading, case reconstruction from seeds."""
import glob
import os

import numpy as np
import torch

from sat_b200 import synth

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def golden_files(prefix):
    return sorted(glob.glob(os.path.join(GOLDEN, prefix + "*.npz")))


def load_fwd_case(path):
    z = np.load(path, allow_pickle=False)
    V, D, B, L, T, tf, ado, att, bert, gate, dropout, seed, H = [int(x) for x in z["cfg"]]
    cfg = dict(V=V, D=D, B=B, L=L, T=T, tf=bool(tf), ado=bool(ado), attention=bool(att), bert=bool(bert),
               gate=bool(gate), dropout=bool(dropout), seed=seed, H=H)
    P = synth.make_params(V, D, H, ado=cfg["ado"], seed=seed, emb_scale=0.5)
    a, cap = synth.make_inputs(B, L, D, T, V, seed=seed, start_token=101 if cfg["bert"] else 0)
    masks = synth.make_dropout_masks(B, T, H) if cfg["dropout"] else None
    return cfg, P, a, cap, masks, z


def load_cap_case(path):
    z = np.load(path, allow_pickle=False)
    V, D, L, beam, ado, att, bert, seed, H, nimg = [int(x) for x in z["cfg"]]
    cfg = dict(V=V, D=D, L=L, beam=beam, ado=bool(ado), attention=bool(att), bert=bool(bert), seed=seed, H=H,
               nimg=nimg, eos_bias=float(z["eos_bias"][0]))
    P = synth.make_params(V, D, H, ado=cfg["ado"], seed=seed, emb_scale=0.5)
    key = "f_out.bias" if cfg["ado"] else "deep_output.bias"
    P[key] = P[key].clone()
    P[key][1] += cfg["eos_bias"]
    a, _ = synth.make_inputs(nimg, L, D, 4, V, seed=seed)
    return cfg, P, a, z


def grad_summary(t):
    f = t.detach().reshape(-1).double().cpu()
    n = f.numel()
    idx = np.unique(np.linspace(0, n - 1, num=min(n, 96)).astype(np.int64))
    return np.concatenate([[f.sum().item(), f.abs().sum().item(), float(n)], f[idx].numpy()])


def check_grad_summary(got, want, name, rtol=2e-4, atol=2e-6):
    assert got[2] == want[2], name
    np.testing.assert_allclose(got[3:], want[3:], rtol=rtol, atol=atol, err_msg=name)
    # abs-sum and sum: scale tolerance by abs-sum
    np.testing.assert_allclose(got[1], want[1], rtol=rtol, atol=atol * want[2], err_msg=name + " abs-sum")
    assert abs(got[0] - want[0]) <= rtol * abs(want[1]) + atol * want[2], name + " sum"


def stub_bert():
    """Offline stand-in for BertModel/BertTokenizer.from_pretrained, same as make_golden.py."""
    import torch.nn as nn
    import transformers

    class _Tok:
        cls_token_id, sep_token_id, pad_token_id = 101, 102, 0

    class _Cfg:
        vocab_size, hidden_size = 30522, 768

    class _Bert:
        config = _Cfg()

        def get_input_embeddings(self):
            return nn.Embedding(30522, 768, padding_idx=0)

    transformers.BertModel.from_pretrained = classmethod(lambda c, n: _Bert())
    transformers.BertTokenizer.from_pretrained = classmethod(lambda c, n: _Tok())
