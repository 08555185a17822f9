"""Training metrics of the reference's train loop (train.py:144-145, 174-181) on the device.

`sequence_accuracy(preds, targets, k, ignore_index)` keeps the signature and the result of utils.py:44-80, but reads the
logits ONCE for any number of k values: the CUDA kernel behind `logit_rank` returns, per (b, t) row, how many logits lie
strictly above the target's; the target is among the top k exactly when that rank is < k.  (The reference runs
`preds.topk(k)` once per k.  On exactly tied logits topk's pick is unspecified; here a tie counts for the target.)
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
    _lib.check(L.sat_logit_rank(preds.data_ptr(), ld, rows, V, targets.data_ptr(), rank.data_ptr(),
                                torch.cuda.current_stream(preds.device).cuda_stream), "sat_logit_rank")
    return rank


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
