"""Kernel timeline of ONE fused train step of the bench workload (c2), from the CUDA activity records kineto collects
(start / end timestamps with the real concurrency, unlike ncu's serialised replays).  Prints every kernel with its start
offset, duration and the idle gap since the previous kernel ended, then per-kernel-name totals and the total idle time.
Usage (GPU box): python tools/timeline.py [--split 0|1] > gpurun_out/timeline.txt"""
import argparse
import collections
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from torch.profiler import ProfilerActivity, profile  # noqa: E402

import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--split", type=int, default=0)
    ap.add_argument("--full", action="store_true", help="print every kernel, not only the totals")
    args = ap.parse_args()
    from sat_b200 import _lib, synth
    from sat_b200.decoder import Decoder
    from sat_b200.trainer import FusedTrainStep
    CFG = bench.CFG
    dev = torch.device("cuda", 0)
    _lib.lib().sat_set_decoder_split(args.split)
    if "SAT_TAIL" in os.environ:                       # 0 = one chain, 1 = two chains, 2 = two chains + register-thin sum over t
        _lib.lib().sat_set_tail_overlap(int(os.environ["SAT_TAIL"]))
    dec = Decoder(CFG["V"], CFG["D"], tf=True, attention=True, compute_dtype=torch.bfloat16)
    dec.load_state_dict(synth.make_params(CFG["V"], CFG["D"], CFG["H"], seed=42))
    dec = dec.to(dev).train()
    tr = FusedTrainStep(dec, lr=1e-4, alpha_c=1.0)
    host = bench.make_host_batches(2, CFG["B"])
    batches = [(a.to(dev), c.to(dev)) for a, c in host]
    for i in range(4):
        tr.step(*batches[i % 2])
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        tr.step(*batches[0])
        tr.step(*batches[1])
        torch.cuda.synchronize()
    evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
    evs.sort(key=lambda e: e.time_range.start)
    # second step only (the first absorbs the profiler's start-up)
    names = [e.name for e in evs]
    adam = [i for i, n in enumerate(names) if "adam_tick" in n.lower()]
    if len(adam) >= 2:
        evs = evs[adam[0] + 1:adam[1] + 1]
    t0 = evs[0].time_range.start
    end_prev = t0
    tot = collections.OrderedDict()
    idle = 0.0
    for e in evs:
        s, d = e.time_range.start - t0, e.time_range.end - e.time_range.start
        gap = e.time_range.start - end_prev
        if gap > 0:
            idle += gap
        end_prev = max(end_prev, e.time_range.end)
        short = e.name.split("(")[0][-70:]
        k = tot.setdefault(short, [0, 0.0, 0.0])
        k[0] += 1; k[1] += d; k[2] += max(gap, 0.0)
        if args.full:
            print(f"{s:10.1f} us  dur {d:8.1f}  gap {gap:7.1f}  {short}")
    span = end_prev - t0
    print(f"# step span {span:.1f} us, idle (no kernel running) {idle:.1f} us, kernels {len(evs)}")
    print(f"# {'count':>5} {'sum us':>10} {'avg us':>8} {'gap-before sum':>14}  kernel")
    for k, (n, d, g) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
        print(f"  {n:5d} {d:10.1f} {d / n:8.2f} {g:14.1f}  {k}")


if __name__ == "__main__":
    main()
