"""Synthetic weights and inputs of the shapes BASELINE.json names.

Weights follow the reference modules' default init *ranges* (nn.Linear / nn.LSTMCell:
U(-1/sqrt(fan_in), +), nn.Embedding: N(0,1)) but are drawn from numpy's legacy
RandomState so they are reproducible across torch versions and machines; the same
dict is loaded into the reference module, the oracle and this package's Decoder.
Keys are the reference's state_dict keys.
"""
from __future__ import annotations

import math

import numpy as np
import torch


def _uniform(rs, shape, bound):
    return torch.from_numpy(rs.uniform(-bound, bound, size=shape).astype(np.float32))


def make_params(V: int, D: int, H: int = 512, ado: bool = False, seed: int = 42,
                emb_scale: float = 1.0) -> dict:
    rs = np.random.RandomState(seed)
    P = {}

    def linear(name, out_f, in_f):
        b = 1.0 / math.sqrt(in_f)
        P[name + ".weight"] = _uniform(rs, (out_f, in_f), b)
        P[name + ".bias"] = _uniform(rs, (out_f,), b)

    P["embedding.weight"] = torch.from_numpy((rs.standard_normal((V, H)) * emb_scale).astype(np.float32))
    linear("init_h", H, D)
    linear("init_c", H, D)
    linear("f_beta", D, H)
    linear("attention.U", H, H)
    linear("attention.W", H, D)
    linear("attention.v", 1, H)
    b = 1.0 / math.sqrt(H)
    P["lstm.weight_ih"] = _uniform(rs, (4 * H, H + D), b)
    P["lstm.weight_hh"] = _uniform(rs, (4 * H, H), b)
    P["lstm.bias_ih"] = _uniform(rs, (4 * H,), b)
    P["lstm.bias_hh"] = _uniform(rs, (4 * H,), b)
    if ado:
        linear("f_h", H, H)
        linear("f_z", H, D)
        linear("f_out", V, H)
    linear("deep_output", V, H)
    return P


def make_inputs(B: int, L: int, D: int, T: int, V: int, seed: int = 42, start_token: int = 0,
                eos: int = 1, pad: int = 3, first_word: int = 4):
    """img_features = relu(randn(B,L,D)) (trunk outputs are post-ReLU); captions int64
    (B, T+1): col 0 = <start>/[CLS], tokens ~ U[first_word, V), one <eos> at a random
    position >= min(5, T) followed by <pad> (mirrors generate_json_data.py:71-78)."""
    rs = np.random.RandomState(seed + 1000)
    a = torch.from_numpy(np.maximum(rs.standard_normal((B, L, D)), 0.0).astype(np.float32))
    cap = rs.randint(first_word, V, size=(B, T + 1)).astype(np.int64)
    cap[:, 0] = start_token
    lo = min(5, T)
    for b in range(B):
        e = rs.randint(lo, T + 1)
        cap[b, e] = eos
        cap[b, e + 1:] = pad
    return a, torch.from_numpy(cap)


def make_dropout_masks(B: int, T: int, H: int, p: float = 0.5, seed: int = 7):
    rs = np.random.RandomState(seed)
    keep = (rs.uniform(size=(B, T, H)) >= p).astype(np.float32) / (1.0 - p)
    return torch.from_numpy(keep)
