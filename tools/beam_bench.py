#!/usr/bin/env python
"""Beam-search throughput of the CUDA path on BASELINE config c4 (secondary metric: beam-5 / beam-10 captions per second).

  python tools/beam_bench.py [--images 128] [--D 512] [--H 512] [--V 2531] [--iters 3] [--cpu-images 2]

Per (dtype, beam): `Decoder.caption_batch` on NI images at once (one process = one GPU; the multi-GPU number is NI images
per GPU, no collective).  Two workloads: "fixed" = end-of-sentence never wins (every image runs all 50 steps and falls
back per Q18 -- the fixed-work number), "eos" = the end token's bias is raised so that captions terminate at realistic
lengths.  With --cpu-images the oracle's per-image beam search (the reference algorithm, `decoder.py:160-269`) is timed on
the host cores on a small sample for the same weights.  Prints one JSON line per case.
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--images", type=int, default=128)
    ap.add_argument("--L", type=int, default=196)
    ap.add_argument("--D", type=int, default=512)
    ap.add_argument("--H", type=int, default=512)
    ap.add_argument("--V", type=int, default=2531)
    ap.add_argument("--iters", type=int, default=3)
    ap.add_argument("--cpu-images", type=int, default=0)
    args = ap.parse_args()
    from sat_b200 import synth
    from sat_b200.decoder import Decoder
    dev = torch.device("cuda", 0)
    if "SAT_ATTN_PART" in os.environ:                       # 0 = automatic, 1 = whole items per CTA, 2 = equal slot counts + combine
        from sat_b200 import _lib
        _lib.lib().sat_set_attn_partition(int(os.environ["SAT_ATTN_PART"]))
    P = synth.make_params(args.V, args.D, args.H, seed=42)
    g = torch.Generator().manual_seed(7)
    a32 = torch.relu(torch.randn(args.images, args.L, args.D, generator=g))
    for workload, eos_bias in (("fixed", -1e4), ("eos", 0.6)):
        Pw = {k: v.clone() for k, v in P.items()}
        Pw["deep_output.bias"][1] += eos_bias                    # token 1 = <eos> (generate_json_data.py:76)
        for dt in (torch.float32, torch.bfloat16):
            dec = Decoder(args.V, args.D, tf=False, ado=False, bert=False, attention=True, compute_dtype=dt)
            dec.load_state_dict(Pw)
            dec = dec.to(dev).eval()
            a = a32.to(dev).to(dt).contiguous()
            for beam in (5, 10):
                dec.caption_batch(a, beam)                        # warm-up (workspace, first-use costs)
                torch.cuda.synchronize()
                times = []
                for _ in range(args.iters):
                    t0 = time.perf_counter()
                    res = dec.caption_batch(a, beam)              # includes the D2H read of the results, as a user sees it
                    torch.cuda.synchronize()
                    times.append(time.perf_counter() - t0)
                best = min(times)
                lens = [len(s) for s, _ in res]
                print(json.dumps({"metric": "beam_captions_per_s", "workload": workload, "beam": beam,
                                  "dtype": "f32" if dt == torch.float32 else "bf16", "images": args.images,
                                  "value": args.images / best, "ms_per_call": best * 1e3,
                                  "mean_caption_len": sum(lens) / len(lens),
                                  "config": f"c4: L={args.L} D={args.D} H={args.H} V={args.V}, <=50 steps"}), flush=True)
        if args.cpu_images > 0:
            from oracle import sat_oracle as O
            torch.set_num_threads(os.cpu_count() or 1)
            for beam in (5, 10):
                t0 = time.perf_counter()
                for i in range(args.cpu_images):
                    O.caption(Pw, a32[i], beam, ado=False, attention=True)
                dt_cpu = time.perf_counter() - t0
                print(json.dumps({"metric": "beam_captions_per_s", "impl": "oracle port on CPU (reference algorithm, fp32)",
                                  "workload": workload, "beam": beam, "images": args.cpu_images,
                                  "cores": os.cpu_count(), "value": args.cpu_images / dt_cpu}), flush=True)


if __name__ == "__main__":
    main()
