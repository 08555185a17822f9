"""CPU oracle for the Show-Attend-and-Tell decoder hot path (TEST INFRASTRUCTURE ONLY).

A plain, functional torch-CPU restatement of the reference's `attention.py` and
`decoder.py` (plus the eight loss lines of `train.py`).  Every function cites the
reference file:line it follows (paths relative to the upstream repo root).

Parity pinning: this restatement is pinned against the *live* reference executed
in the build container (`tests/golden/make_golden.py` imports the upstream
modules from /root/reference, runs them on seeded weights/inputs and commits the
outputs under tests/golden/*.npz).  `tests/test_oracle_golden.py` checks the
restatement against those vectors on CPU.  The reference itself ships no tests or
golden vectors, so the live-reference fixtures are the pin.

The product package never imports this module; the CUDA path must fail loudly
when its extension is missing.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Sequence, Tuple

import torch

Params = Dict[str, torch.Tensor]


# ----------------------------------------------------------------------------
# primitive layers
# ----------------------------------------------------------------------------
# Optional emulation of the CUDA path's STORAGE precision (bf16 mode): every tensor the kernels keep in bf16 -- GEMM
# operands (activations and packed weights), Wa, P1 / P2 / comb of the advanced deep output -- is rounded to bf16 at
# the same points, with a straight-through gradient; all arithmetic stays fp32.  The reference itself has no such mode
# (it is fp32 only); this exists because a bf16 run cannot be compared tightly with fp32 THROUGH ReLUs: a pre-activation
# within rounding noise of zero flips its mask and changes a whole gradient row.  With the rounding points aligned the masks
# agree and the comparison measures the kernels, not the quantisation.  Used by tests only (`with O.emulate_storage(...)`).
_STORAGE_DTYPE = None


class emulate_storage:
    def __init__(self, dtype):
        self.dtype = dtype

    def __enter__(self):
        global _STORAGE_DTYPE
        self.prev, _STORAGE_DTYPE = _STORAGE_DTYPE, self.dtype
        return self

    def __exit__(self, *exc):
        global _STORAGE_DTYPE
        _STORAGE_DTYPE = self.prev


def _q(x: torch.Tensor) -> torch.Tensor:
    if _STORAGE_DTYPE is None or _STORAGE_DTYPE == torch.float32:
        return x
    return x + (x.detach().to(_STORAGE_DTYPE).to(x.dtype) - x.detach())


def _linear(P: Params, name: str, x: torch.Tensor, quant: bool = True) -> torch.Tensor:
    """nn.Linear: y = x W^T + b."""
    if quant:
        return _q(x) @ _q(P[name + ".weight"]).t() + P[name + ".bias"]
    return x @ P[name + ".weight"].t() + P[name + ".bias"]


def attention_forward(P: Params, a: torch.Tensor, h: torch.Tensor):
    """attention.py:14-21.  NOTE the naming swap: `attention.U` acts on
    the hidden state, `attention.W` on the image features."""
    U_h = _linear(P, "attention.U", h).unsqueeze(1)            # attention.py:15
    W_s = _q(_linear(P, "attention.W", a))                     # attention.py:16
    att = torch.tanh(W_s + U_h)                                # attention.py:17
    e = _linear(P, "attention.v", att, quant=False).squeeze(2)  # attention.py:18 (fp32 in the kernels)
    alpha = torch.softmax(e, dim=1)                            # attention.py:19
    context = (a * alpha.unsqueeze(2)).sum(1)                  # attention.py:20
    return context, alpha


def init_lstm_state(P: Params, a: torch.Tensor):
    """decoder.py:137-147."""
    avg = a.mean(dim=1)
    c = torch.tanh(_linear(P, "init_c", avg))
    h = torch.tanh(_linear(P, "init_h", avg))
    return h, c


def lstm_cell(P: Params, x: torch.Tensor, h: torch.Tensor, c: torch.Tensor):
    """nn.LSTMCell used at decoder.py:115 / :197; gate order i,f,g,o."""
    g = _q(x) @ _q(P["lstm.weight_ih"]).t() + P["lstm.bias_ih"] + _q(h) @ _q(P["lstm.weight_hh"]).t() + P["lstm.bias_hh"]
    i, f, gg, o = g.chunk(4, dim=1)
    c2 = torch.sigmoid(f) * c + torch.sigmoid(i) * torch.tanh(gg)
    h2 = torch.sigmoid(o) * torch.tanh(c2)
    return h2, c2


def advanced_deep_output(P: Params, h, context, emb, out_mask=None):
    """decoder.py:149-158 (un-gated context, final ReLU on the logits: SURVEY Q8).  `out_mask` (tests only): use this 0/1
    mask in place of the final ReLU's own (logit > 0) -- the teacher-style check for the ReLU, see decoder_forward."""
    h_t = _q(torch.relu(_linear(P, "f_h", h)))
    z_t = _q(torch.relu(_linear(P, "f_z", context)))
    pre = _linear(P, "f_out", h_t + z_t + _q(emb))
    return torch.relu(pre) if out_mask is None else pre * out_mask


def _step(P: Params, a, h, c, emb, *, attention: bool, gate: bool = True):
    """One decode step up to and including the LSTM cell (decoder.py:97-115 / :185-197)."""
    B, L, _ = a.shape
    if attention:
        context, alpha = attention_forward(P, a, h)                        # decoder.py:98
        if gate:
            g = torch.sigmoid(_linear(P, "f_beta", h))                      # decoder.py:99
            gated = g * context                                             # decoder.py:100
        else:                       # build extension: gating scalar disabled
            gated = context
    else:
        alpha = torch.full((B, L), 1.0 / L, dtype=a.dtype, device=a.device)                # decoder.py:103
        context = a.mean(dim=1)                                             # decoder.py:104
        gated = context                                                     # decoder.py:105
    x = torch.cat((emb, gated), dim=1)                                      # decoder.py:109/112
    h, c = lstm_cell(P, x, h, c)                                            # decoder.py:115
    return h, c, context, alpha


def decoder_forward(P: Params, a: torch.Tensor, captions: torch.Tensor, *, tf: bool, ado: bool,
                    attention: bool, gate: bool = True, start_token: int = 0,
                    dropout_masks: Optional[torch.Tensor] = None, feed_tokens: Optional[torch.Tensor] = None,
                    relu_masks: Optional[torch.Tensor] = None):
    """decoder.py:69-135.  `dropout_masks` (B,T,H) holds the *scaled* keep mask
    (0 or 1/(1-p)) applied to h before the output layer (decoder.py:121-125);
    None means eval mode (identity).  Tensors are created on `a.device`, so the same restatement runs on the GPU
    for the full-size parity tests.  `feed_tokens` (B,T), free-running mode only: feed these ids as step inputs
    instead of this function's own argmax (decoder.py:132) -- the teacher-style check of SURVEY 7.2: a bf16 run's
    near-tie argmax may legitimately differ from the fp32 one, after which the two trajectories are incomparable.
    `relu_masks` (B,T,V) bool, advanced deep output only: the same idea for the final ReLU on the V logits (decoder.py:158) --
    the run under test says which logits it saw as positive.  A logit within bf16 noise of zero flips its mask; with V logits
    per row a fraction f of flipped masks moves a gradient by ~sqrt(f) in relative L2 (0.3 % of flips -> 5 %), which says
    nothing about the kernels."""
    B = a.shape[0]
    h, c = init_lstm_state(P, a)                                            # decoder.py:74
    T = captions.shape[1] - 1                                               # decoder.py:77
    E = P["embedding.weight"]
    if tf:
        cap_emb = E[captions]                                               # decoder.py:87
    else:
        prev = E[torch.full((B,), start_token, dtype=torch.long, device=a.device)]   # decoder.py:79-82,89
    V = P["deep_output.weight"].shape[0]
    preds = torch.zeros(B, T, V, dtype=a.dtype, device=a.device)            # decoder.py:92
    alphas = torch.zeros(B, T, a.shape[1], dtype=a.dtype, device=a.device)  # decoder.py:93
    for t in range(T):                                                      # decoder.py:96
        emb = cap_emb[:, t] if tf else prev
        h, c, context, alpha = _step(P, a, h, c, emb, attention=attention, gate=gate)
        hd = h if dropout_masks is None else h * dropout_masks[:, t]
        if ado:
            out = advanced_deep_output(P, hd, context, emb,                 # decoder.py:118-123
                                       None if relu_masks is None else relu_masks[:, t].to(a.dtype))
        else:
            out = _linear(P, "deep_output", hd)                             # decoder.py:125
        preds[:, t] = out                                                   # decoder.py:127
        alphas[:, t] = alpha                                                # decoder.py:128
        if not tf:
            prev = E[out.max(1)[1] if feed_tokens is None or t + 1 >= T else feed_tokens[:, t + 1]]   # decoder.py:132-133
    return preds, alphas


def train_loss(preds: torch.Tensor, alphas: torch.Tensor, captions: torch.Tensor, alpha_c: float = 1.0):
    """train.py:135,150-154,161-162.  pack_padded_sequence with all lengths = T-1 is the
    time-major flattening of [:, :T-1]; CE is an unmasked mean over B*(T-1)
    rows; the doubly-stochastic regulariser sums alphas over TIME."""
    targets = captions[:, 1:]
    T = preds.shape[1]
    pp = preds[:, :T - 1].transpose(0, 1).reshape(-1, preds.shape[2])
    pt = targets[:, :T - 1].transpose(0, 1).reshape(-1)
    ce = torch.nn.functional.cross_entropy(pp, pt)
    reg = alpha_c * ((1.0 - alphas.sum(1)) ** 2).mean()
    return ce + reg


# ----------------------------------------------------------------------------
# beam search
# ----------------------------------------------------------------------------
def caption(P: Params, a1: torch.Tensor, beam_size: int, *, ado: bool, attention: bool,
            bert: bool = False, start_token: int = 0, gate: bool = True, max_steps: int = 50):
    """decoder.py:160-269, restated for ONE image `a1` of shape (L, D).

    Returns (sentence, alpha, fallback) where fallback=True mirrors the reference's
    "No completed sentences found" branch (decoder.py:256-258): sentence == [0] and
    alpha is the last step's (beam_live, L) tensor.  Otherwise sentence is list[int] and
    alpha list[list[float]] whose first row is all ones (decoder.py:173)."""
    L = a1.shape[0]
    a = a1.unsqueeze(0).expand(beam_size, L, a1.shape[1])                   # generate_caption.py:89
    E = P["embedding.weight"]
    V = (P["f_out.weight"] if ado else P["deep_output.weight"]).shape[0]
    prev_words = torch.full((beam_size,), start_token, dtype=torch.long, device=a1.device)    # decoder.py:166-169
    sentences = prev_words.unsqueeze(1)
    top_preds = torch.zeros(beam_size, 1, dtype=a1.dtype, device=a1.device)
    alphas = torch.ones(beam_size, 1, L, dtype=a1.dtype, device=a1.device)
    done_s: List[List[int]] = []
    done_a: List[List[List[float]]] = []
    done_p: List[float] = []
    step = 1
    h, c = init_lstm_state(P, a)                                            # decoder.py:180
    alpha = None
    while True:
        emb = E[prev_words]                                                 # decoder.py:183
        h, c, context, alpha = _step(P, a, h, c, emb, attention=attention, gate=gate)
        if ado:
            out = advanced_deep_output(P, h, context, emb)                  # decoder.py:199-201
        else:
            out = _linear(P, "deep_output", h)                              # decoder.py:203
        out = top_preds.expand_as(out) + out                                # decoder.py:204 (raw logits, Q13)
        if step == 1:
            top_preds, top_words = out[0].topk(beam_size, 0, True, True)    # decoder.py:207 (Q14)
        else:
            top_preds, top_words = out.reshape(-1).topk(beam_size, 0, True, True)
        prev_idx = (top_words // V).long()                                  # decoder.py:210
        next_idx = (top_words % V).long()                                   # decoder.py:211
        sentences = torch.cat((sentences[prev_idx], next_idx.unsqueeze(1)), dim=1)       # decoder.py:220
        alphas = torch.cat((alphas[prev_idx], alpha[prev_idx].unsqueeze(1)), dim=1)      # decoder.py:221
        stop = (1, 0) if bert else (1, 102)                                 # decoder.py:224-229 (Q15)
        incomplete = [i for i, w in enumerate(next_idx.tolist()) if w not in stop]
        complete = sorted(set(range(len(next_idx))) - set(incomplete))
        if complete:                                                        # decoder.py:234-237
            done_s.extend(sentences[complete].tolist())
            done_a.extend(alphas[complete].tolist())
            done_p.extend(top_preds[complete].tolist())
        beam_size -= len(complete)                                          # decoder.py:238
        if beam_size == 0:
            break
        sentences = sentences[incomplete]                                   # decoder.py:244-250 (Q16)
        alphas = alphas[incomplete]
        h = h[prev_idx[incomplete]]
        c = c[prev_idx[incomplete]]
        a = a[prev_idx[incomplete]]
        top_preds = top_preds[incomplete].unsqueeze(1)
        prev_words = next_idx[incomplete]
        if step > max_steps:                                                # decoder.py:252-253 (Q17)
            break
        step += 1
    if not done_p:                                                          # decoder.py:256-258 (Q18)
        return [0], alpha, True
    idx = done_p.index(max(done_p))                                         # decoder.py:265 (first max)
    return done_s[idx], done_a[idx], False


def sequence_accuracy(preds: torch.Tensor, targets: torch.Tensor, k: int, ignore_index: int = 0) -> float:
    """utils.py:44-80 restated: accuracy at k in percent over the tokens that are not `ignore_index`.
    preds (B, T, V), targets (B, T); train.py:144-145 calls it with targets = captions[:, 1:] and k = 1, 5."""
    _, topk = preds.topk(k, dim=2, largest=True, sorted=True)                  # utils.py:55
    mask = targets.ne(ignore_index)                                            # utils.py:61
    correct = topk.eq(targets.unsqueeze(-1).expand_as(topk)) * mask.unsqueeze(-1)   # utils.py:64-65
    correct_total = correct.any(dim=2).float().sum()                           # utils.py:68-71
    n = mask.sum()                                                             # utils.py:77
    return (correct_total.item() * 100.0 / n.item()) if n.item() > 0 else 0    # utils.py:80
