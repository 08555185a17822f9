"""Multi-GPU beam search (BASELINE config c4): the images are sharded over the ranks (one process per GPU),
every rank runs `Decoder.caption_batch` on ITS shard -- beam search is embarrassingly parallel per image, so the data path
has no collective -- and the variable-length results are gathered on one rank as padded tensors.

The reference captions one image per call on the CPU (generate_caption.py:88-90) and has no multi-device code.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def shard_range(n_items: int, rank: int, world: int):
    """Contiguous, near-equal shard [lo, hi) of n_items for `rank` (the first n_items % world ranks get one more)."""
    base, rem = divmod(int(n_items), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def gather_caption_results(raw: dict, group=None, dst: int = 0, with_alphas: bool = True):
    """Gather the padded per-image outputs of `Decoder.caption_batch_raw` from every rank on `dst`.
    raw: lens (n,), status (n,), live (n,) int32; tokens (n, maxlen) int64; alphas (n, maxlen, L), last_alpha (n, beam, L) fp32.
    Shards may differ in size.  Returns the concatenated dict on dst (rank order = image order), None elsewhere."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return raw
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    dev = raw["tokens"].device
    n = torch.tensor([raw["tokens"].shape[0]], device=dev, dtype=torch.int64)
    counts = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(counts, n, group=group)
    counts = [int(c.item()) for c in counts]
    nmax = max(counts)
    keys = ["lens", "status", "live", "tokens"] + (["alphas", "last_alpha"] if with_alphas else [])
    out = {}
    gdst = dist.get_global_rank(group, dst) if group is not None else dst
    for k in keys:
        t = raw[k]
        if t.shape[0] < nmax:               # pad the shard to the largest one: gather wants equal shapes
            t = torch.cat((t, t.new_zeros((nmax - t.shape[0],) + tuple(t.shape[1:]))), 0)
        t = t.contiguous()
        bufs = [torch.empty_like(t) for _ in range(world)] if rank == dst else None
        dist.gather(t, bufs, dst=gdst, group=group)
        if rank == dst:
            out[k] = torch.cat([b[:c] for b, c in zip(bufs, counts)], 0)
    return out if rank == dst else None


def caption_sharded(decoder, img_features, beam_size, *, group=None, dst: int = 0, max_steps: int = 50, with_alphas: bool = True):
    """`img_features` (n_local, L, D) is THIS rank's shard of the images.  Returns on `dst` the list of (sentence, alpha)
    pairs of all images in rank order, with the reference's per-image return convention (decoder.py:256-269); None on the
    other ranks.  `with_alphas=False` gathers only the sentences (alpha = None for completed captions)."""
    raw = decoder.caption_batch_raw(img_features, beam_size, max_steps=max_steps)
    allr = gather_caption_results(raw, group=group, dst=dst, with_alphas=with_alphas)
    if allr is None:
        return None
    return decoder.results_from_raw(allr)
