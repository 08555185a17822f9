"""Runs the UNMODIFIED reference modules staged in oracle/_ref (see build_ref.py) on the host cores.

TEST INFRASTRUCTURE ONLY -- the timed CPU baseline of bench.py (`--impl reference`, `cpu_baseline`) and a checker for
tests/.  Nothing under show-attend-and-tell_b200/ imports this.

What is the reference's and what is restated here:
  * `Decoder(...)`, `Decoder.forward`, `Decoder.caption`, `Attention.forward`: upstream code, imported from oracle/_ref,
    untouched.  The one injection is the module global `decoder.mps_device` (upstream defines it on Apple-Silicon only,
    decoder.py:5-6 -- SURVEY Q1).
  * the loss: train.py cannot be imported here (matplotlib / nltk / skimage / wandb are not installed), so its eight loss
    lines (train.py:135,150-154,161-162) are restated in `train_loss` below, pack_padded_sequence call included.
  * optimizer: torch.optim.Adam(decoder.parameters(), lr) as train.py:71.
"""
from __future__ import annotations

import importlib
import os
import sys
import time

import torch
from torch.nn.utils.rnn import pack_padded_sequence

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")


def available() -> bool:
    return all(os.path.exists(os.path.join(REF_DIR, f)) for f in ("attention.py", "decoder.py"))


_MOD = None


def ref_decoder_module(device="cpu"):
    """The staged upstream `decoder` module with `mps_device` injected."""
    global _MOD
    if _MOD is None:
        if not available():
            raise RuntimeError("oracle/_ref is not staged; run `python oracle/build_ref.py` where /root/reference exists")
        # upstream files are top-level modules named `attention` / `decoder` (decoder.py: `from attention import Attention`)
        sys.path.insert(0, REF_DIR)
        try:
            for name in ("attention", "decoder"):
                if name in sys.modules and not getattr(sys.modules[name], "__file__", "").startswith(REF_DIR):
                    del sys.modules[name]
            _MOD = importlib.import_module("decoder")
        finally:
            sys.path.remove(REF_DIR)
    _MOD.mps_device = torch.device(device)
    return _MOD


def train_loss(preds, alphas, captions, alpha_c, cross_entropy_loss):
    """train.py:135,150-154,161-162, restated line for line."""
    targets = captions[:, 1:]                                                                   # :135
    packed_targets = pack_padded_sequence(targets, [len(tar) - 1 for tar in targets], batch_first=True)[0]   # :150
    packed_preds = pack_padded_sequence(preds, [len(pred) - 1 for pred in preds], batch_first=True)[0]       # :151
    att_regularization = alpha_c * ((1 - alphas.sum(1)) ** 2).mean()                             # :154
    loss = cross_entropy_loss(packed_preds, packed_targets)                                      # :161
    loss += att_regularization                                                                   # :162
    return loss


def make_decoder(params, V, D, *, tf=True, ado=False, attention=True, device="cpu"):
    """Upstream Decoder with the given state_dict (reference key names)."""
    mod = ref_decoder_module(device)
    dec = mod.Decoder(V, D, tf=tf, ado=ado, bert=False, attention=attention)
    dec.load_state_dict(params)
    return dec.to(device)


def train_step_rate(params, a, captions, *, steps, warmup, threads, tf=True, ado=False, attention=True, lr=1e-4,
                    alpha_c=1.0):
    """Times `steps` train steps (after `warmup`) of the upstream Decoder on the host: forward (train mode, dropout on),
    train.py loss, backward, Adam.  Returns (captions/s, mean ms per step, last loss)."""
    torch.set_num_threads(threads)
    dec = make_decoder(params, params["deep_output.weight"].shape[0], a.shape[2], tf=tf, ado=ado, attention=attention)
    dec.train()
    opt = torch.optim.Adam(dec.parameters(), lr=lr)                  # train.py:71
    ce = torch.nn.CrossEntropyLoss()                                 # train.py:73
    times, loss = [], None
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        opt.zero_grad()                                              # train.py:133
        preds, alphas = dec(a, captions)                             # train.py:134
        loss = train_loss(preds, alphas, captions, alpha_c, ce)
        loss.backward()                                              # train.py:163
        opt.step()                                                   # train.py:164
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    ms = 1e3 * sum(times) / len(times)
    return a.shape[0] / (ms / 1e3), ms, float(loss)


def caption_rate(params, feats, beam, *, threads, ado=False, attention=True):
    """Upstream `Decoder.caption` looped over the images of `feats` (n, L, D) as generate_caption.py:89-90 does
    (features expanded to beam rows).  Returns (captions/s, list of (sentence, alpha))."""
    import contextlib
    import io
    torch.set_num_threads(threads)
    dec = make_decoder(params, params["deep_output.weight"].shape[0], feats.shape[2], tf=False, ado=ado, attention=attention)
    dec.eval()
    out = []
    t0 = time.perf_counter()
    with torch.no_grad(), contextlib.redirect_stdout(io.StringIO()):
        for i in range(feats.shape[0]):
            f = feats[i:i + 1].expand(beam, feats.shape[1], feats.shape[2])
            out.append(dec.caption(f, beam))
    dt = time.perf_counter() - t0
    return feats.shape[0] / dt, out
