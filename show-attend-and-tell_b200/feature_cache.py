"""HBM-resident cache of encoder features.

The reference recomputes `img_features = encoder(imgs)` for every batch of every epoch (train.py:132), although the
VGG19 trunk is frozen (encoder.py:29-31) and the images are fixed tensors preloaded by the dataset (dataset.py:36-41):
the features are constants.  Kept in HBM once -- Flickr8k at L=196 x D=512 in bf16 is 1.6 GB of a B200's 180 GB -- a train
step's host traffic shrinks from the feature batch (205 MB at c2) to the batch's indices and captions (~45 KB), and
the 4.45 s (CPU) encoder pass per batch disappears.

    cache = FeatureCache(n_images, L, D, dtype=torch.bfloat16, device="cuda")
    cache.put(indices, encoder(imgs))            # once per image; accepts the encoder's permuted view (Q20)
    feats = cache.gather(batch_indices, out=buf)  # (B, L, D) contiguous, one HBM-bound copy kernel (sat_gather_items)

The gather is a plain byte copy (bit-exact); an index outside the cache zero-fills its row and sets a device flag that
`check()` turns into the IndexError an eager tensor index would have raised.
"""
from __future__ import annotations

import torch

from . import _lib


class FeatureCache:
    def __init__(self, n_items: int, L: int, D: int, dtype=torch.bfloat16, device="cuda"):
        dev = torch.device(device)
        if dev.type != "cuda":
            raise RuntimeError("sat_b200: FeatureCache lives in GPU memory (the hot path has no CPU fallback)")
        if dtype not in (torch.float32, torch.bfloat16):
            raise TypeError("FeatureCache supports float32 and bfloat16")
        self.n_items, self.L, self.D, self.dtype = int(n_items), int(L), int(D), dtype
        self.store = torch.zeros(self.n_items, self.L, self.D, device=dev, dtype=dtype)
        self.item_bytes = self.L * self.D * self.store.element_size()
        if self.item_bytes % 16:
            raise ValueError("L * D * itemsize must be a multiple of 16 bytes")
        self.filled = torch.zeros(self.n_items, device=dev, dtype=torch.bool)
        self._err = torch.zeros(1, device=dev, dtype=torch.int32)

    @property
    def device(self):
        return self.store.device

    def nbytes(self) -> int:
        return self.store.numel() * self.store.element_size()

    def put(self, indices, features):
        """store[indices[i]] = features[i]  (features (n, L, D), any float dtype / strides; cast to the cache dtype)."""
        idx = torch.as_tensor(indices, dtype=torch.int64, device=self.device)
        if features.shape[1:] != (self.L, self.D) or features.shape[0] != idx.numel():
            raise RuntimeError(f"features {tuple(features.shape)} do not match ({idx.numel()}, {self.L}, {self.D})")
        self.store.index_copy_(0, idx, features.detach().to(device=self.device, dtype=self.dtype))
        self.filled[idx] = True

    def gather(self, indices, out=None):
        """(B,) int64 indices (device, or host -> copied) -> contiguous (B, L, D) features in the cache dtype."""
        idx = indices if torch.is_tensor(indices) else torch.as_tensor(indices)
        idx = idx.to(device=self.device, dtype=torch.int64, non_blocking=True).contiguous()
        n = idx.numel()
        if out is None:
            out = torch.empty(n, self.L, self.D, device=self.device, dtype=self.dtype)
        elif out.shape != (n, self.L, self.D) or out.dtype != self.dtype or not out.is_contiguous() or out.device != self.device:
            raise RuntimeError("out must be a contiguous (B, L, D) tensor of the cache dtype on the cache device")
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().sat_gather_items(self.store.data_ptr(), self.item_bytes, self.n_items, idx.data_ptr(), n,
                                                   out.data_ptr(), self._err.data_ptr(),
                                                   torch.cuda.current_stream(self.device).cuda_stream), "sat_gather_items")
        return out

    def check(self):
        """Raise IndexError if any gather since the last check saw an index outside [0, n_items) (host sync)."""
        if int(self._err.item()):
            self._err.zero_()
            raise IndexError(f"FeatureCache.gather: index out of range [0, {self.n_items})")
