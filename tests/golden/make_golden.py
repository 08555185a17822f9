"""Generate golden vectors by running the UNMODIFIED upstream reference modules live.

Run in the build container only (needs /root/reference):
    python tests/golden/make_golden.py
Writes tests/golden/*.npz.  Weights/inputs are regenerated from seeds by
`sat_b200.synth` (numpy RandomState), so fixtures hold only configs + outputs.
The only touches to the reference are the ones SURVEY.md section 0.1 lists as unavoidable
off-Mac: inject `decoder.mps_device` (Q1), stub BERT `from_pretrained` (Q23), and for
the dropout / no-gate cases swap the `dropout` / `sigmoid` sub-modules (Q5, Q9).
"""
import io
import contextlib
import os
import sys

import numpy as np
import torch
import torch.nn as nn

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference")

import decoder as refdec  # noqa: E402  (upstream decoder.py)
from sat_b200 import synth  # noqa: E402

refdec.mps_device = torch.device("cpu")  # Q1
torch.set_num_threads(4)


class FixedDropout(nn.Module):
    """Applies masks[:, t] on the t-th call (Q9)."""

    def __init__(self, masks):
        super().__init__()
        self.masks, self.t = masks, 0

    def forward(self, x):
        y = x * self.masks[:, self.t]
        self.t += 1
        return y


class Ones(nn.Module):
    def forward(self, x):
        return torch.ones_like(x)


def stub_bert():
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


def build_ref(V, D, tf, ado, bert, attention, P):
    dec = refdec.Decoder(V, D, tf=tf, ado=ado, bert=bert, attention=attention)
    missing = dec.load_state_dict(P, strict=True)
    dec.eval()
    return dec


def ref_loss(preds, alphas, captions, alpha_c=1.0):
    """train.py:135,150-154,161-162 verbatim semantics (pack_padded_sequence form)."""
    from torch.nn.utils.rnn import pack_padded_sequence
    targets = captions[:, 1:]
    packed_targets = pack_padded_sequence(targets, [len(t) - 1 for t in targets], batch_first=True)[0]
    packed_preds = pack_padded_sequence(preds, [len(p) - 1 for p in preds], batch_first=True)[0]
    reg = alpha_c * ((1 - alphas.sum(1)) ** 2).mean()
    return nn.CrossEntropyLoss()(packed_preds, packed_targets) + reg


def grad_summary(t):
    f = t.detach().reshape(-1).double()
    n = f.numel()
    idx = np.unique(np.linspace(0, n - 1, num=min(n, 96)).astype(np.int64))
    return np.concatenate([[f.sum().item(), f.abs().sum().item(), float(n)], f[idx].numpy()])


def fwd_case(name, *, V, D, B, L, T, tf, ado, attention, bert=False, gate=True, dropout=False,
             seed=42, backward=True, sample_preds=1):
    H = 768 if bert else 512
    P = synth.make_params(V, D, H, ado=ado, seed=seed, emb_scale=0.5)
    a, cap = synth.make_inputs(B, L, D, T, V, seed=seed, start_token=101 if bert else 0)
    dec = build_ref(V, D, tf, ado, bert, attention, P)
    if not gate:
        dec.sigmoid = Ones()
    if dropout:
        masks = synth.make_dropout_masks(B, T, H)
        dec.dropout = FixedDropout(masks)
    preds, alphas = dec(a, cap)
    out = dict(cfg=np.array([V, D, B, L, T, int(tf), int(ado), int(attention), int(bert), int(gate),
                             int(dropout), seed, H], dtype=np.int64),
               preds=preds.detach().numpy()[:, :, ::sample_preds].copy(),
               preds_argmax=preds.detach().argmax(2).numpy(),
               alphas=alphas.detach().numpy())
    if backward:
        loss = ref_loss(preds, alphas, cap)
        loss.backward()
        out["loss"] = np.array([loss.item()])
        names = []
        for n, p in dec.named_parameters():
            if p.grad is None:
                names.append(n + ":none")
                continue
            names.append(n)
            out["grad." + n] = grad_summary(p.grad)
        out["grad_names"] = np.array(names)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print("wrote", name, {k: v.shape for k, v in out.items() if hasattr(v, "shape")} if False else "")


def caption_case(name, *, V, D, L, beam, ado, attention, eos_bias, bert=False, seed=42, nimg=3):
    H = 768 if bert else 512
    P = synth.make_params(V, D, H, ado=ado, seed=seed, emb_scale=0.5)
    key = "f_out.bias" if ado else "deep_output.bias"
    P[key] = P[key].clone()
    P[key][1] += eos_bias  # random-init never emits <eos> without a nudge
    a, _ = synth.make_inputs(nimg, L, D, 4, V, seed=seed)
    dec = build_ref(V, D, True, ado, bert, attention, P)
    out = dict(cfg=np.array([V, D, L, beam, int(ado), int(attention), int(bert), seed, H, nimg], dtype=np.int64),
               eos_bias=np.array([eos_bias], dtype=np.float64))
    with torch.no_grad():
        for i in range(nimg):
            feats = a[i:i + 1].expand(beam, L, D)       # generate_caption.py:89
            buf = io.StringIO()
            with contextlib.redirect_stdout(buf):
                sent, alpha = dec.caption(feats, beam)
            fallback = "No completed sentences found" in buf.getvalue()
            out[f"sent{i}"] = np.array(sent, dtype=np.int64)
            out[f"alpha{i}"] = (alpha.numpy() if torch.is_tensor(alpha) else np.array(alpha, dtype=np.float32))
            out[f"fallback{i}"] = np.array([int(fallback)])
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print("wrote", name, [out[f"sent{i}"].tolist() for i in range(nimg)])


def main():
    small = dict(V=37, D=24, B=3, L=10, T=4)
    for att in (0, 1):
        for ado in (0, 1):
            for tf in (0, 1):
                fwd_case(f"fwd_att{att}_ado{ado}_tf{tf}", tf=bool(tf), ado=bool(ado), attention=bool(att), **small)
    fwd_case("fwd_dropout", tf=True, ado=False, attention=True, dropout=True, **small)
    fwd_case("fwd_dropout_ado", tf=True, ado=True, attention=True, dropout=True, **small)
    fwd_case("fwd_nogate", tf=True, ado=False, attention=True, gate=False, **small)
    fwd_case("fwd_c1like", V=2531, D=512, B=2, L=196, T=3, tf=True, ado=False, attention=True, sample_preds=5)
    stub_bert()
    fwd_case("fwd_bert", V=30522, D=16, B=2, L=6, T=3, tf=False, ado=False, attention=True, bert=True,
             gate=True, sample_preds=61)
    for beam, bias in ((3, 0.0), (5, 0.0), (5, 0.1), (10, 0.0), (10, 0.2), (5, -50.0)):
        caption_case(f"cap_beam{beam}_b{int(bias * 10)}".replace("-", "m"), V=37, D=24, L=10, beam=beam,
                     ado=False, attention=True, eos_bias=bias)
    caption_case("cap_ado_beam5", V=37, D=24, L=10, beam=5, ado=True, attention=True, eos_bias=1.0)
    caption_case("cap_noatt_beam5", V=37, D=24, L=10, beam=5, ado=False, attention=False, eos_bias=0.1)
    caption_case("cap_v2531_beam5_b0", V=2531, D=512, L=196, beam=5, ado=False, attention=True, eos_bias=0.0, nimg=2)
    caption_case("cap_v2531_beam5_b15", V=2531, D=512, L=196, beam=5, ado=False, attention=True, eos_bias=0.15, nimg=2)


if __name__ == "__main__":
    main()
