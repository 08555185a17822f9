"""Training metrics of the reference's train loop (train.py:144-145, 174-181) on the device.

`sequence_accuracy(preds, targets, k, ignore_index)` keeps the signature and the result of utils.py:44-80, but reads the
logits ONCE for any number of k values: the CUDA kernel behind `logit_rank` returns, per (b, t) row, the position of the
target's logit in the row's descending order; the target is among the top k exactly when that rank is < k.  (The
reference runs `preds.topk(k)` once per k.)  Ties are broken towards the lower vocabulary index, so a block of tied
logits admits at most k entries, as topk does -- with the advanced deep output the final ReLU (decoder.py:158) makes
blocks of exact zeros the norm, and counting every tie for the target would score an untrained model as correct.  A NaN
target logit is never correct.

`caption_token_count` is utils.py:101-107 (`calculate_caption_lengths`, train.py:174-177) without the `.item()`.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib


def logit_rank(preds: torch.Tensor, targets: torch.Tensor, *, ld: int | None = None) -> torch.Tensor:
    """preds (..., V) fp32 CUDA logits (last dim contiguous, rows `ld` apart), targets (...) int64 -> int32 ranks (...)."""
    _lib.require_cuda(preds, "preds")
    if preds.dtype != torch.float32:
        raise TypeError("preds must be float32 (decoder.py:92)")
    V = preds.shape[-1]
    if ld is None:
        preds = preds.contiguous()
        ld = V
    targets = targets.to(device=preds.device, dtype=torch.int64).contiguous()
    rows = targets.numel()
    if preds.numel() // V != rows:
        raise RuntimeError(f"preds {tuple(preds.shape)} and targets {tuple(targets.shape)} do not match")
    rank = torch.empty(targets.shape, device=preds.device, dtype=torch.int32)
    L = _lib.lib()
    with torch.cuda.device(preds.device):
        _lib.check(L.sat_logit_rank(preds.data_ptr(), ld, rows, V, targets.data_ptr(), rank.data_ptr(),
                                    torch.cuda.current_stream(preds.device).cuda_stream), "sat_logit_rank")
    return rank


def caption_token_count(captions: torch.Tensor, skip_tokens) -> torch.Tensor:
    """utils.py:101-107: number of ids in `captions` (B, Tcap) that are none of `skip_tokens` (at most three ids: pad, start,
    eos -- train.py:174-177).  Returns a device int64 tensor of shape (1,); the reference's `.item()` is left to the caller."""
    _lib.require_cuda(captions, "captions")
    skip = [int(x) for x in (skip_tokens.tolist() if torch.is_tensor(skip_tokens) else skip_tokens)]
    if len(skip) > 3:
        raise ValueError("caption_token_count supports up to three skip tokens (pad, start, eos)")
    skip += [-1] * (3 - len(skip))
    cap = captions.to(torch.int64)
    if cap.dim() != 2 or cap.stride(1) != 1:
        cap = cap.reshape(cap.shape[0], -1).contiguous()
    out = torch.empty(1, device=cap.device, dtype=torch.int64)
    with torch.cuda.device(cap.device):
        _lib.check(_lib.lib().sat_count_tokens(cap.data_ptr(), cap.shape[0], cap.shape[1], cap.stride(0), skip[0], skip[1], skip[2],
                                               out.data_ptr(), torch.cuda.current_stream(cap.device).cuda_stream), "sat_count_tokens")
    return out


def calculate_caption_lengths(captions, skip_tokens):
    """Drop-in for utils.py:101-107 (returns a Python int like the reference's `.item()`)."""
    return int(caption_token_count(captions, skip_tokens).item())


def greedy_tokens(preds: torch.Tensor) -> torch.Tensor:
    """train.py:272,297 `torch.max(preds, dim=2)[1]`: (B, T, V) fp32 logits -> (B, T) int64 ids of the first maximum."""
    _lib.require_cuda(preds, "preds")
    if preds.dtype != torch.float32:
        raise TypeError("preds must be float32 (decoder.py:92)")
    preds = preds.contiguous()
    V = preds.shape[-1]
    out = torch.empty(preds.shape[:-1], device=preds.device, dtype=torch.int64)
    with torch.cuda.device(preds.device):
        _lib.check(_lib.lib().sat_argmax_rows(preds.data_ptr(), V, out.numel(), V, out.data_ptr(),
                                              torch.cuda.current_stream(preds.device).cuda_stream), "sat_argmax_rows")
    return out


def accuracy_counts(rank: torch.Tensor, targets: torch.Tensor, ks=(1, 5), ignore_index: int = 0) -> torch.Tensor:
    """Device tensor [correct@k for k in ks ..., non-padding tokens] (float32); no host sync."""
    mask = targets.to(rank.device).ne(ignore_index)
    out = [((rank < k) & mask).sum() for k in ks] + [mask.sum()]
    return torch.stack(out).to(torch.float32)


def sequence_accuracy(preds, targets, k, ignore_index=0, tokenizer=None):
    """Drop-in for utils.py:44-80: accuracy at k in percent over the non-padding tokens (a Python float, like the
    reference's `.item()` arithmetic)."""
    c = accuracy_counts(logit_rank(preds, targets), targets, (k,), ignore_index).tolist()
    return (c[0] * 100.0 / c[1]) if c[1] > 0 else 0
