#!/usr/bin/env python
"""Fused train step (forward + train.py loss + backward + fused Adam) of the CUDA path at the full size of every training
config BASELINE.json names, one GPU each (bench.py times c2 only; these are the other rows of SURVEY 8(d)):

  c2        B=256 D=2048 H=512 V=2531  T=20 teacher forcing
  c3-D512   B=128 D=512  H=512 V=10000 T=50 teacher forcing
  c3-D2048  B=128 D=2048 H=512 V=10000 T=50 teacher forcing
  c5        B=128 D=2048 H=768 V=30522 T=20 free running (tf=False), BERT-sized frozen embedding (offline stub of
            BertModel.from_pretrained), context gate off

bf16, synthetic inputs resident in HBM, random-init weights, train-mode dropout.  Prints one JSON line per config with the
algorithmic HBM bytes of the attention steps (SURVEY 8(d) with this design's backward: T*[2*L*(H+D)*s + L*H*s] + T*L*H*s per
caption) and the fraction of the measured copy peak the WHOLE step reaches on that figure.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

CONFIGS = {
    "c2": dict(B=256, D=2048, V=2531, T=20, tf=True, bert=False, gate=True),
    "c3-D512": dict(B=128, D=512, V=10000, T=50, tf=True, bert=False, gate=True),
    "c3-D2048": dict(B=128, D=2048, V=10000, T=50, tf=True, bert=False, gate=True),
    "c5": dict(B=128, D=2048, V=30522, T=20, tf=False, bert=True, gate=False),
}


def stub_bert():
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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="c2,c3-D512,c3-D2048,c5")
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=6, help=">= 6: each of the two batch buffers needs three steps before its CUDA graph exists")
    ap.add_argument("--ado", type=int, default=0, help="1 = advanced deep output (the flag set train_models.py passes)")
    ap.add_argument("--graph", type=int, default=1, help="1 = replay the step as one CUDA graph (what bench.py does)")
    args = ap.parse_args()
    from sat_b200.decoder import Decoder
    from sat_b200.trainer import FusedTrainStep
    import bench
    peak, peak_src = bench.read_peaks()
    dev = torch.device("cuda", 0)
    L = 196
    for name in args.configs.split(","):
        c = CONFIGS[name]
        torch.manual_seed(42)
        if c["bert"]:
            stub_bert()
        dec = Decoder(c["V"], c["D"], tf=c["tf"], ado=bool(args.ado), bert=c["bert"], attention=True, compute_dtype=torch.bfloat16,
                      gate=c["gate"]).to(dev).train()
        H, V, B, T, D = dec.embedding_size, dec.vocabulary_size, c["B"], c["T"], c["D"]
        tr = FusedTrainStep(dec, lr=1e-4, alpha_c=1.0, distributed=False, use_cuda_graph=bool(args.graph))
        batches = []
        for i in range(2):
            g = torch.Generator().manual_seed(42 + i)
            a = torch.relu(torch.randn(B, L, D, generator=g)).to(torch.bfloat16).to(dev)
            cap = torch.randint(4, V, (B, T + 1), generator=g)
            cap[:, 0] = dec.start_token
            batches.append((a, cap.to(dev)))
        for i in range(args.warmup):
            loss = tr.step(*batches[i % 2])
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(args.steps):
            loss = tr.step(*batches[i % 2])
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / args.steps
        s = 2
        # per decode step: a + Wa read by the forward and by the backward kernel, the backward's score gradients written; once per
        # train step: the rebuild of sum_t dpre (Wa read, the sum written) -- the same accounting as bench.py's step_hbm_floor_ms
        attn_bytes = B * T * (2 * L * (H + D) * s + L * 4) + 2 * B * L * H * s
        floor_ms = attn_bytes / (peak * 1e9) * 1e3
        lv = [float(x) for x in loss.tolist()]
        print(json.dumps({"config": name, "shape": {"B": B, "L": L, "D": D, "H": H, "V": V, "T": T, "tf": c["tf"], "gate": c["gate"],
                                                    "frozen_embedding": c["bert"], "ado": bool(args.ado)},
                          "launch": "cuda graph" if tr.graph_replays else "eager",
                          "dtype": "bf16", "ms_per_step": ms, "captions_per_s": B / (ms / 1e3),
                          "attention_algorithmic_bytes": attn_bytes, "hbm_floor_ms": floor_ms, "step_frac_of_hbm_floor": floor_ms / ms,
                          "peak_gbs": peak, "peak_source": peak_src, "loss": lv[2], "finite": all(map(lambda x: x == x and abs(x) < 1e30, lv)),
                          "workspace_GB": tr._ws.numel() / 1e9}), flush=True)
        del tr, dec, batches
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
