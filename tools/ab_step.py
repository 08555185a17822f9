"""A/B timing of the fused c2 train step under library tuning knobs (device-resident, CUDA events, median of repeats).
  python tools/ab_step.py "pf=0,0" "pf=32,32" "pf=64,64;part=2" ...
Knobs per variant (';'-separated): pf=F,B (sat_set_attn_prefetch), part=M (sat_set_attn_partition), graph=0|1 (CUDA-graph replay), wide=0|1 (sat_set_gemm_wide_tile),
ado=0|1, tail=0..3 (sat_set_tail_overlap), dpre=0|1 (sat_set_dpre_recompute), buckets=2|3 (FusedTrainStep.buckets), l2=F,B[,PIN_MB] (sat_set_attn_l2_policy: Wa hint fwd, bwd, pinned feature megabytes).  Prints one line per variant: ms/step (median of 5 x 20 steps), attention fwd / bwd event-bracket averages."""
import ctypes
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402


def main():
    from sat_b200 import _lib, synth
    from sat_b200.decoder import Decoder
    from sat_b200.trainer import FusedTrainStep
    L = _lib.lib()
    CFG = bench.CFG
    dev = torch.device("cuda", 0)
    host = bench.make_host_batches(2, CFG["B"])
    batches = [(a.to(dev), c.to(dev)) for a, c in host]
    variants = sys.argv[1:] or ["pf=0,0"]
    trainers = {}
    for v in variants:
        kn = dict(kv.split("=") for kv in v.split(";") if kv)
        pf = [int(x) for x in kn.get("pf", "0,8").split(",")]
        L.sat_set_attn_prefetch(pf[0], pf[1])
        L.sat_set_attn_partition(int(kn.get("part", 0)))
        L.sat_set_gemm_wide_tile(int(kn.get("wide", 1)))
        l2 = [int(x) for x in kn.get("l2", "1,0,0").split(",")] + [0]
        L.sat_set_attn_l2_policy(l2[0], l2[1], ctypes.c_int64(l2[2] * 1000000))
        if "tail" in kn:
            L.sat_set_tail_overlap(int(kn["tail"]))
        graph, ado = int(kn.get("graph", 0)), int(kn.get("ado", 0))
        key = (graph, ado)
        if key not in trainers:
            dec = Decoder(CFG["V"], CFG["D"], tf=True, ado=bool(ado), attention=True, compute_dtype=torch.bfloat16)
            dec.load_state_dict(synth.make_params(CFG["V"], CFG["D"], CFG["H"], ado=bool(ado), seed=42))
            trainers[key] = FusedTrainStep(dec.to(dev).train(), lr=1e-4, use_cuda_graph=bool(graph))
        tr = trainers[key]
        L.sat_set_dpre_recompute(int(kn.get("dpre", 1)))
        tr.buckets = int(kn.get("buckets", 2))
        tr._graphs.clear(); tr._seen.clear()          # knobs are baked into captured graphs
        tr._ws_key = None                              # ... and into the workspace layout
        for i in range(8):
            tr.step(*batches[i % 2])
        torch.cuda.synchronize()
        reps = []
        for _ in range(5):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(20):
                tr.step(*batches[i % 2])
            e1.record()
            torch.cuda.synchronize()
            reps.append(e0.elapsed_time(e1) / 20)
        ms = (ctypes.c_float * 3)()
        cnt = (ctypes.c_int * 3)()
        if not graph:
            L.sat_profile(1)
            for i in range(2):
                tr.step(*batches[i % 2])
            torch.cuda.synchronize()
            L.sat_profile_read(ms, cnt)
            L.sat_profile(0)
        f = ms[0] / max(cnt[0], 1) * 1e3
        b = ms[1] / max(cnt[1], 1) * 1e3
        print(f"{v:32s} ms/step median {statistics.median(reps):.4f}  min {min(reps):.4f}  max {max(reps):.4f}   attn fwd {f:.1f} us  bwd {b:.1f} us  loss {tr.loss[2].item():.4f}", flush=True)


if __name__ == "__main__":
    main()
