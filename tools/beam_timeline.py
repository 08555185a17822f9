"""Kernel timeline of one batched beam-search call (c4 shape) from kineto's CUDA activity records: per-kernel-name totals,
host gaps, and the per-decode-step period.  Usage (GPU box): python tools/beam_timeline.py [--images 128] [--beam 5]"""
import argparse
import collections
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from torch.profiler import ProfilerActivity, profile  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--images", type=int, default=128)
    ap.add_argument("--beam", type=int, default=5)
    ap.add_argument("--V", type=int, default=2531)
    ap.add_argument("--D", type=int, default=512)
    ap.add_argument("--full", action="store_true")
    args = ap.parse_args()
    from sat_b200 import synth
    from sat_b200.decoder import Decoder
    dev = torch.device("cuda", 0)
    P = synth.make_params(args.V, args.D, 512, seed=42)
    P["deep_output.bias"] = P["deep_output.bias"].clone()
    P["deep_output.bias"][1] -= 1e4                 # <eos> never wins: all 51 iterations run (fixed work)
    g = torch.Generator().manual_seed(7)
    a32 = torch.relu(torch.randn(args.images, 196, args.D, generator=g))
    for dt in (torch.float32, torch.bfloat16):
        dec = Decoder(args.V, args.D, tf=False, attention=True, compute_dtype=dt)
        dec.load_state_dict(P)
        dec = dec.to(dev).eval()
        a = a32.to(dev).to(dt).contiguous()
        dec.caption_batch(a, args.beam)
        torch.cuda.synchronize()
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            dec.caption_batch(a, args.beam)
            torch.cuda.synchronize()
        evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
        evs.sort(key=lambda e: e.time_range.start)
        t0 = evs[0].time_range.start
        end_prev, idle = t0, 0.0
        tot = collections.OrderedDict()
        for e in evs:
            d = e.time_range.end - e.time_range.start
            gap = e.time_range.start - end_prev
            if gap > 0:
                idle += gap
            end_prev = max(end_prev, e.time_range.end)
            short = e.name.split("(")[0][-80:]
            k = tot.setdefault(short, [0, 0.0, 0.0])
            k[0] += 1; k[1] += d; k[2] += max(gap, 0.0)
            if args.full:
                print(f"{e.time_range.start - t0:10.1f} us  dur {d:8.1f}  gap {gap:7.1f}  {short}")
        span = end_prev - t0
        print(f"# dtype {dt} images {args.images} beam {args.beam}: span {span:.1f} us, idle {idle:.1f} us, kernels {len(evs)}, "
              f"per decode step {span / 51:.1f} us")
        print(f"# {'count':>5} {'sum us':>10} {'avg us':>8} {'gap-before sum':>14}  kernel")
        for k, (n, d, gp) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
            print(f"  {n:5d} {d:10.1f} {d / n:8.2f} {gp:14.1f}  {k}")


if __name__ == "__main__":
    main()
