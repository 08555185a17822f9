#!/usr/bin/env python
"""Benchmark of the decoder hot path on BASELINE.json config c2 (see DESIGN.md "Measurement").

  python bench.py --gpus N --steps K --warmup W                 our arm (one process per GPU under torchrun for N>1)
  python bench.py --impl reference --gpus N --steps K --warmup W  the reference's own modules on the host cores

A "step" is one teacher-forced train step of the decoder: forward + train.py loss + backward (+ NCCL gradient
all-reduce for N>1) + Adam, batch 256 per GPU (weak scaling), synthetic ResNet152-shaped features
(L=196, D=2048), H=512, V=2531, T=20, bf16 compute.  Prints ONE JSON line on rank 0.

Keys beyond the base contract: `e2e` (features uploaded from pinned host memory every step), `e2e_cached` (features gathered
from the HBM-resident FeatureCache: the step uploads indices + captions), `roofline` (dominant attention kernel, CUDA-event
brackets on its stream), `cpu_baseline` (the reference modules of oracle/_ref timed on this box's host cores, N=1 only),
`beam` (BASELINE's secondary metric, config c4: 128 images per GPU, beam 5 and 10, every rank captions its shard).
"""
import argparse
import ctypes
import json
import os
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

CFG = dict(B=256, L=196, D=2048, H=512, V=2531, T=20)
WORKLOAD = ("c2: teacher-forced bf16 train step of the soft-attention LSTM decoder, ResNet152-shaped features "
            "L=196 x D=2048, H=512, V=2531, T=20, batch 256 per GPU")
METRIC, UNIT = "train_captions_per_s", "captions/s"
CPU_SAMPLE_B = 32            # captions per CPU step of the reference arm (a bounded sample of the 256-caption c2 batch)


_JSON_OUT = None


def read_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def bind_to_gpu_numa_node(local_rank):
    """Pin this process (and the pinned host buffers it is about to first-touch) to the NUMA node its GPU hangs off.  With 8
    ranks each uploading 205 MB per step, pages that all live on one socket -- or a rank running across the inter-socket
    link -- cap the host side long before PCIe does (r1: 55 GB/s per rank alone, 26 GB/s at N=8).  Best effort: returns a
    description, never raises."""
    try:
        props = torch.cuda.get_device_properties(local_rank)
        bus = "%04x:%02x:%02x.0" % (getattr(props, "pci_domain_id", 0), props.pci_bus_id, props.pci_device_id)
        with open(f"/sys/bus/pci/devices/{bus}/numa_node") as f:
            node = int(f.read().strip())
        if node < 0:
            return {"numa_node": None, "note": "no NUMA affinity reported for the GPU"}
        with open(f"/sys/devices/system/node/node{node}/cpulist") as f:
            spec = f.read().strip()
        cpus = set()
        for part in spec.split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        allowed = os.sched_getaffinity(0)
        cpus &= allowed
        if not cpus:
            return {"numa_node": node, "note": "node CPUs not in this process's cpuset"}
        os.sched_setaffinity(0, cpus)
        return {"numa_node": node, "cpus": len(cpus)}
    except Exception as e:                                       # noqa: BLE001
        return {"numa_node": None, "note": f"{type(e).__name__}"}


class ClockSampler:
    """SM clock + throttle reasons sampled DURING the timed regions.  NVML is polled from a thread every few ms (the timed
    region is only tens of ms long, far below nvidia-smi's start-up time); `mark(True/False)` brackets the regions that count."""
    REASONS = (("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20), ("sw_power_cap", 0x4))

    def __init__(self, dev):
        self.rows, self.thread, self.handle, self.nv = [], None, None, None
        self.stop_flag, self.active, self.err = False, False, None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            try:
                uuid = str(torch.cuda.get_device_properties(dev).uuid)
                if not uuid.startswith("GPU-"):
                    uuid = "GPU-" + uuid
                self.handle = pynvml.nvmlDeviceGetHandleByUUID(uuid)
            except Exception:
                vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
                idx = dev.index or 0
                if vis and all(x.strip().isdigit() for x in vis.split(",")):
                    idx = int(vis.split(",")[idx])
                self.handle = pynvml.nvmlDeviceGetHandleByIndex(idx)
            self.sm_max = float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
        except Exception as e:                                   # noqa: BLE001
            self.err = f"nvml unavailable: {type(e).__name__}"
            self.handle = None

    def _poll(self):
        nv, h = self.nv, self.handle
        while not self.stop_flag:
            try:
                sm = float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                try:
                    rs = int(nv.nvmlDeviceGetCurrentClocksEventReasons(h))
                except Exception:
                    rs = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(h))
                self.rows.append((self.active, sm, rs))
            except Exception as e:                               # noqa: BLE001
                self.err = f"nvml query failed: {type(e).__name__}"
                return
            time.sleep(0.002)

    def start(self):
        if self.handle is not None:
            self.thread = threading.Thread(target=self._poll, daemon=True)
            self.thread.start()

    def mark(self, on):
        self.active = on

    def stop(self):
        if self.thread is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [self.err or "nvml unavailable"], "samples": 0}
        self.stop_flag = True
        self.thread.join(timeout=2)
        rows = [r for r in self.rows if r[0]] or self.rows
        sm = [r[1] for r in rows]
        bits = 0
        for r in rows:
            bits |= r[2]
        reasons = [n for n, b in self.REASONS if bits & b]
        if self.err:
            reasons.append(self.err)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": self.sm_max, "reasons": reasons,
                "samples": len(sm), "sampled": "NVML polled every ~2 ms inside the timed regions (value + e2e + e2e_cached)"}


def make_host_batches(n, B, seed0=42):
    """n distinct synthetic batches in pinned host memory (bf16 features + int64 captions)."""
    from sat_b200 import synth
    out = []
    for i in range(n):
        g = torch.Generator().manual_seed(seed0 + i)
        a = torch.relu(torch.randn(B, CFG["L"], CFG["D"], generator=g)).to(torch.bfloat16).pin_memory()
        _, cap = synth.make_inputs(B, 1, 8, CFG["T"], CFG["V"], seed=seed0 + i)
        out.append((a, cap.pin_memory()))
    return out


# ---------------------------------------------------------------------------------------------------------------------------
# the reference arm: the reference's OWN attention.py / decoder.py (staged in oracle/_ref by oracle/build_ref.py) on the host cores
# ---------------------------------------------------------------------------------------------------------------------------
def cpu_reference_rate(sample_B, steps, warmup, threads):
    """One train step of the c2 workload on a `sample_B`-caption sample: upstream Decoder.forward (train mode) + the train.py
    loss + backward + torch.optim.Adam, fp32, `threads` host threads.  Falls back to the oracle port (kind "port") only if
    oracle/_ref was not staged.  Returns (captions/s, ms per step, kind, description)."""
    from sat_b200 import synth
    P = synth.make_params(CFG["V"], CFG["D"], CFG["H"], seed=42, emb_scale=1.0)
    a, cap = synth.make_inputs(sample_B, CFG["L"], CFG["D"], CFG["T"], CFG["V"], seed=42)
    from oracle import ref_runner
    if ref_runner.available():
        val, ms, _ = ref_runner.train_step_rate(P, a, cap, steps=steps, warmup=warmup, threads=threads)
        kind = "reference"
        what = "the reference's own attention.py + decoder.py (oracle/_ref, unmodified; mps_device injected) + train.py's loss lines restated"
    else:
        from oracle import sat_oracle as O
        torch.set_num_threads(threads)
        Pg = {k: v.clone().requires_grad_(True) for k, v in P.items()}
        masks = synth.make_dropout_masks(sample_B, CFG["T"], CFG["H"])
        opt = torch.optim.Adam(list(Pg.values()), lr=1e-4)
        times = []
        for i in range(warmup + steps):
            t0 = time.perf_counter()
            opt.zero_grad(set_to_none=True)
            preds, alphas = O.decoder_forward(Pg, a, cap, tf=True, ado=False, attention=True, dropout_masks=masks)
            O.train_loss(preds, alphas, cap, 1.0).backward()
            opt.step()
            if i >= warmup:
                times.append(time.perf_counter() - t0)
        ms = 1e3 * sum(times) / len(times)
        val, kind, what = sample_B / (ms / 1e3), "port", "oracle port of attention.py / decoder.py / train.py loss (oracle/_ref not staged)"
    sample = (f"{what}; CPU, torch {torch.__version__}, fp32, {threads} threads; each step = one full train step (forward, loss, "
              f"backward, Adam) on a {sample_B}-caption sample of the c2 batch (same L, D, H, V, T); {steps} timed steps after "
              f"{warmup} warm-up ({ms:.0f} ms/step)")
    return val, ms, kind, sample


def run_reference(args, rank):
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    warm = max(1, args.warmup)
    val, ms, kind, sample = cpu_reference_rate(CPU_SAMPLE_B, args.steps, warm, threads)
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": warm, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": {"workload": WORKLOAD, "sample_batch": CPU_SAMPLE_B},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": kind, "sample": sample},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def count_my_launches(fn):
    """Kernel launches of ONE step, counted with the CUDA profiler (kineto); all kernels in a step are ours or NCCL's.
    The same activity records give each attention kernel's EXPOSED time: from the later of (its own start, the end of the
    kernel before it) to its end.  With programmatic dependent launch a kernel starts while its predecessor drains, so this
    is what one launch adds to the step -- the CUDA-event brackets of the roofline section cannot overlap like that."""
    try:
        from torch.profiler import ProfilerActivity, profile
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            fn()
            torch.cuda.synchronize()
        evs = [ev for ev in prof.events() if ev.device_type == torch.autograd.DeviceType.CUDA
               and "memcpy" not in ev.name.lower() and "memset" not in ev.name.lower()]
        evs.sort(key=lambda e: e.time_range.start)
        exposed, prev_end = {}, None
        for ev in evs:
            st, en = ev.time_range.start, ev.time_range.end
            for key, tag in (("attn_step_fwd", "fwd_stream_kernel"), ("attn_step_bwd", "bwd_stream_kernel")):
                if tag in ev.name and prev_end is not None:
                    exposed.setdefault(key, []).append(en - max(st, prev_end))
            prev_end = en if prev_end is None else max(prev_end, en)
        return len(evs), {k: sum(v) / len(v) / 1e3 for k, v in exposed.items()}
    except Exception:
        return None, {}


# ---------------------------------------------------------------------------------------------------------------------------
# secondary metric: beam-search captions/s (config c4), every rank captions its own shard of the images
# ---------------------------------------------------------------------------------------------------------------------------
def beam_block(dev, rank, world, barrier, max_over_ranks, images=128, V=2531, D=512, H=512, Lloc=196, steps=51):
    """Decoder.caption_batch (beam 5 and 10, <= 51 iterations, fp32 = the bit-exact mode, and bf16) on `images` VGG-shaped
    feature maps per GPU resident in HBM; the end token never wins, so every image runs all 51 iterations (fixed work).
    Each call ends with the gather of the sentences on rank 0 (sharded.caption_sharded); rate = all ranks' images / max time.
    roofline: the attention step's algorithmic bytes per image per decode step, L (H + D) s, independent of the beam width
    (beam rows share the image)."""
    from sat_b200 import synth
    from sat_b200.decoder import Decoder
    from sat_b200 import sharded
    P = synth.make_params(V, D, H, seed=42)
    P["deep_output.bias"] = P["deep_output.bias"].clone()
    P["deep_output.bias"][1] -= 1e4                          # token 1 = <eos> (generate_json_data.py:76)
    g = torch.Generator().manual_seed(7 + rank)
    a32 = torch.relu(torch.randn(images, Lloc, D, generator=g))
    peak, _ = read_peaks()
    out = {"workload": f"c4: {images} images per GPU ({world * images} in all), L={Lloc} D={D} H={H} V={V}, beam 5 / 10, "
                       f"{steps} iterations (no early end), sentences gathered on rank 0", "unit": "captions/s",
           "sharding": f"{world} rank(s) x {images} images, no data-path collective"}
    for dt, key, es in ((torch.float32, "f32", 4), (torch.bfloat16, "bf16", 2)):
        dec = Decoder(V, D, tf=False, ado=False, bert=False, attention=True, compute_dtype=dt)
        dec.load_state_dict(P)
        dec = dec.to(dev).eval()
        a = a32.to(dev).to(dt).contiguous()
        for beam in (5, 10):
            sharded.caption_sharded(dec, a, beam, with_alphas=False)
            barrier()
            best = None
            for _ in range(3):
                barrier()
                t0 = time.perf_counter()
                res = sharded.caption_sharded(dec, a, beam, with_alphas=False)
                torch.cuda.synchronize()
                dtm = max_over_ranks((time.perf_counter() - t0) * 1e3)
                best = dtm if best is None else min(best, dtm)
            if rank == 0:
                assert len(res) == world * images
            attn_bytes = images * steps * Lloc * (H + D) * es
            out[f"{key}_beam{beam}"] = {"value": world * images / (best / 1e3), "ms_per_call": best,
                                        "us_per_decode_step": best * 1e3 / steps,
                                        "roofline": {"bound": "hbm", "kernel": "attn_step_fwd (beam rows share the image)",
                                                     "algorithmic_bytes_per_call": attn_bytes,
                                                     "achieved": attn_bytes / (best / 1e3) / 1e9, "peak": peak, "unit": "GB/s",
                                                     "frac": attn_bytes / (best / 1e3) / 1e9 / peak,
                                                     "note": "whole call (GEMMs, top-k, bookkeeping included) against the attention bytes; "
                                                             "the 51 MB (bf16) / 103 MB (fp32) of a + Wa are L2-resident across steps"}}
    return out


def run_ours(args, rank, world, local_rank):
    import torch.distributed as dist
    numa = bind_to_gpu_numa_node(local_rank)               # before any pinned allocation (first touch)
    from sat_b200 import _lib, synth
    from sat_b200.decoder import Decoder
    from sat_b200.feature_cache import FeatureCache
    from sat_b200.trainer import FusedTrainStep
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    L = _lib.lib()
    if "SAT_DEC_SPLIT" in os.environ:                      # A/B switches of library knobs (see include/sat_b200.h)
        L.sat_set_decoder_split(int(os.environ["SAT_DEC_SPLIT"]))
    if "SAT_FUSE_CELL" in os.environ:
        L.sat_set_fused_cell_backward(int(os.environ["SAT_FUSE_CELL"]))
    if "SAT_TAIL" in os.environ:
        L.sat_set_tail_overlap(int(os.environ["SAT_TAIL"]))
    if "SAT_STATIC_W" in os.environ:
        L.sat_set_gemm_static_weights(int(os.environ["SAT_STATIC_W"]))
    if "SAT_ATTN_PF" in os.environ:                        # "fwd_slots,bwd_slots"
        L.sat_set_attn_prefetch(*[int(x) for x in os.environ["SAT_ATTN_PF"].split(",")])
    if "SAT_ATTN_PART" in os.environ:
        L.sat_set_attn_partition(int(os.environ["SAT_ATTN_PART"]))
    use_graph = os.environ.get("SAT_CUDA_GRAPH", "1") != "0"
    B, T = CFG["B"], CFG["T"]
    dec = Decoder(CFG["V"], CFG["D"], tf=True, ado=False, bert=False, attention=True, compute_dtype=torch.bfloat16)
    dec.load_state_dict(synth.make_params(CFG["V"], CFG["D"], CFG["H"], seed=42))
    dec = dec.to(dev).train()
    # N > 1: gradients cross NVLink as bf16 (half the bytes of the one exchange step); SAT_GRAD_EXCHANGE=f32 switches it off
    gx = torch.float32 if (world == 1 or os.environ.get("SAT_GRAD_EXCHANGE", "bf16") == "f32") else torch.bfloat16
    trainer = FusedTrainStep(dec, lr=1e-4, alpha_c=1.0, use_cuda_graph=use_graph, grad_exchange_dtype=gx)
    if "SAT_BUCKETS" in os.environ:                                               # experiment knob: 2 / 3 = bucketed exchange under the backward tail
        trainer.buckets, trainer.overlap_allreduce = int(os.environ["SAT_BUCKETS"]), True
    if "SAT_OVERLAP" in os.environ:                                               # experiment knob: 1 = early bucket(s) under the backward tail
        trainer.overlap_allreduce = os.environ["SAT_OVERLAP"] == "1"
    host = make_host_batches(2, B, seed0=42 + 10 * rank)
    dev_batches = [(a.to(dev, non_blocking=True), c.to(dev, non_blocking=True)) for a, c in host]
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return t.item()
        return ms

    graph_note = None
    try:
        for i in range(max(args.warmup, 6)):               # >= 3 warm-up steps per buffer pair: the third captures its CUDA graph
            trainer.step(*dev_batches[i % 2])
        torch.cuda.synchronize()
    except Exception as e:                                 # noqa: BLE001  (capture refused: run the same kernels launch by launch)
        if not use_graph:
            raise
        graph_note = f"CUDA-graph capture failed ({type(e).__name__}: {str(e)[:120]}); eager launches"
        torch.cuda.synchronize()
        trainer.use_cuda_graph = False
        trainer._graphs.clear()
        for i in range(args.warmup):
            trainer.step(*dev_batches[i % 2])
        torch.cuda.synchronize()

    # ---- device-resident throughput ("value"): K steps, repeated; the median repeat is reported ---------------------------
    sampler = ClockSampler(dev)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = []
    for _ in range(args.repeats):
        barrier()
        sampler.mark(True)
        e0.record()
        for i in range(args.steps):
            loss = trainer.step(*dev_batches[i % 2])
        e1.record()
        barrier()
        sampler.mark(False)
        reps.append(max_over_ranks(e0.elapsed_time(e1)) / args.steps)
    ms_step = statistics.median(reps)
    value = world * B / (ms_step / 1e3)
    loss_val = loss[2].item()
    trainer.check_status()

    # ---- end to end through the public API with host buffers ("e2e") ------------------------------
    copy_stream = torch.cuda.Stream(dev)
    bufs = [(torch.empty_like(dev_batches[0][0]), torch.empty_like(dev_batches[0][1])) for _ in range(2)]
    ready = [torch.cuda.Event(), torch.cuda.Event()]
    freed = [torch.cuda.Event(), torch.cuda.Event()]
    loss_host = torch.zeros(4).pin_memory()

    def e2e_loop(n, prefetch):
        cur = torch.cuda.current_stream(dev)
        for s in range(2):
            freed[s].record(cur)
        prefetch(0)
        for i in range(n):
            s = i % 2
            if i + 1 < n:
                prefetch(i + 1)
            cur.wait_event(ready[s])
            l = trainer.step(bufs[s][0], bufs[s][1])
            freed[s].record(cur)
            loss_host.copy_(l, non_blocking=True)
            cur.synchronize()                                  # the step's loss is on the host (train.py:179)
        trainer.check_status(loss_host)
        return float(loss_host[2])

    def prefetch_features(i):
        s = i % 2
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(freed[s])
            bufs[s][0].copy_(host[s][0], non_blocking=True)
            bufs[s][1].copy_(host[s][1], non_blocking=True)
            ready[s].record(copy_stream)

    def timed_e2e(prefetch):
        e2e_loop(6, prefetch)                                  # warm-up (captures the graphs of the two staging buffers)
        rr = []
        for _ in range(args.repeats):
            barrier()
            sampler.mark(True)
            e0.record()
            e2e_loop(args.steps, prefetch)
            e1.record()
            barrier()
            sampler.mark(False)
            rr.append(max_over_ranks(e0.elapsed_time(e1)) / args.steps)
        return statistics.median(rr), rr

    # SAT_BENCH_SKIP_SIDE=1 (tools/r2_call9.sh: the ncu launch list of this command) shortens the side measurements so that the
    # profiled process is mostly train steps; the JSON line keeps its shape
    skip_side = os.environ.get("SAT_BENCH_SKIP_SIDE", "0") == "1"
    e2e_ms, e2e_reps = timed_e2e(prefetch_features)
    e2e_val = world * B / (e2e_ms / 1e3)
    h2d = host[0][0].numel() * host[0][0].element_size() + host[0][1].numel() * 8
    d2h = 16

    # ---- end to end with the HBM-resident feature cache ("e2e_cached"): a step uploads indices + captions -----------------
    n_cache = 4 * B                                            # 4 batches of distinct features = 822 MB of the 180 GB
    cache = FeatureCache(n_cache, CFG["L"], CFG["D"], dtype=torch.bfloat16, device=dev)
    for j in range(4):
        cache.put(torch.arange(j * B, (j + 1) * B), dev_batches[j % 2][0] if j < 2 else dev_batches[j % 2][0].flip(0))
    gidx = torch.Generator().manual_seed(1234 + rank)
    host_idx = [torch.randperm(n_cache, generator=gidx)[:B].pin_memory() for _ in range(2)]
    idx_dev = [torch.empty(B, device=dev, dtype=torch.int64) for _ in range(2)]

    def prefetch_cached(i):
        s = i % 2
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(freed[s])
            idx_dev[s].copy_(host_idx[s], non_blocking=True)
            bufs[s][1].copy_(host[s][1], non_blocking=True)
            cache.gather(idx_dev[s], out=bufs[s][0])           # HBM -> HBM, on the copy stream, under the previous step
            ready[s].record(copy_stream)

    e2ec_ms, e2ec_reps = timed_e2e(prefetch_cached)
    cache.check()
    e2ec_val = world * B / (e2ec_ms / 1e3)
    h2d_cached = B * 8 + host[0][1].numel() * 8
    clocks = sampler.stop() if rank == 0 else None

    # ---- kernel-level roofline (CUDA events around every attention-step launch, on its own stream; eager launches) -------
    graph_was = trainer.use_cuda_graph
    trainer.use_cuda_graph = False
    ms = (ctypes.c_float * 3)()
    cnt = (ctypes.c_int * 3)()
    trainer.step(*dev_batches[0])
    torch.cuda.synchronize()
    L.sat_profile(1)
    nprof = 2
    for i in range(nprof):
        trainer.step(*dev_batches[i % 2])
    torch.cuda.synchronize()
    _lib.check(L.sat_profile_read(ms, cnt))
    L.sat_profile(0)
    s = 2  # bf16
    Lc, D, H = CFG["L"], CFG["D"], CFG["H"]
    bytes_fwd = B * Lc * (H + D) * s
    # a, Wa read once + this step's score gradients de (B x L fp32) written once: sum_t dpre is rebuilt after the loop in one
    # pass over Wa (misc::dpre_sum, library default sat_set_dpre_recompute(1)); round 1 wrote a (B, L, H) dpre slab per step
    bytes_bwd = B * (Lc * (H + D) * s + Lc * 4)
    bytes_dpre_sum = 2 * B * Lc * H * s                  # the rebuild: Wa read, sum_t dpre written (once per train step)
    peak, peak_src = read_peaks()
    kern = {}
    for name, idx, nbytes in (("attn_step_fwd", 0, bytes_fwd), ("attn_step_bwd", 1, bytes_bwd)):
        if cnt[idx] > 0:
            avg_ms = ms[idx] / cnt[idx]
            kern[name] = {"avg_ms": avg_ms, "launches_per_step": cnt[idx] // nprof, "ms_per_step": ms[idx] / nprof,
                          "algorithmic_bytes": nbytes, "achieved_gbs": nbytes / avg_ms / 1e6}
    kern["gemm_all"] = {"ms_per_step": ms[2] / nprof, "launches_per_step": cnt[2] // nprof}
    dom = max(("attn_step_fwd", "attn_step_bwd"), key=lambda k: kern.get(k, {}).get("ms_per_step", 0.0))
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            traffic = json.load(f).get(dom)
    except Exception:
        pass
    roofline = {"bound": "hbm", "kernel": dom, "achieved": kern[dom]["achieved_gbs"], "peak": peak, "unit": "GB/s",
                "frac": kern[dom]["achieved_gbs"] / peak, "traffic": traffic, "peak_source": peak_src,
                "share_of_step": kern[dom]["ms_per_step"] / ms_step, "kernels": kern,
                "step_hbm_floor_ms": (T * (bytes_fwd + bytes_bwd) + bytes_dpre_sum) / peak / 1e6,
                "step_frac_of_hbm_floor": (T * (bytes_fwd + bytes_bwd) + bytes_dpre_sum) / peak / 1e6 / ms_step}

    launches, exposed = count_my_launches(lambda: trainer.step(*dev_batches[0]))
    trainer.use_cuda_graph = graph_was
    for name, ems in exposed.items():
        if name in kern:
            kern[name]["exposed_avg_ms"] = ems
            kern[name]["exposed_gbs"] = kern[name]["algorithmic_bytes"] / ems / 1e6
    # host -> device bandwidth of the pinned feature batch alone (what bounds e2e when it exceeds the device step)
    h2d_ev0, h2d_ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    h2d_ev0.record()
    for _ in range(3):
        bufs[0][0].copy_(host[0][0], non_blocking=True)
    h2d_ev1.record()
    torch.cuda.synchronize()
    h2d_ms = max_over_ranks(h2d_ev0.elapsed_time(h2d_ev1) / 3)
    del cache, bufs
    torch.cuda.empty_cache()
    # ---- secondary metric of BASELINE.json: beam-search captions/s (config c4), all ranks -----------------------------
    beam = None
    try:
        beam = {"skipped": "SAT_BENCH_SKIP_SIDE"} if skip_side else beam_block(dev, rank, world, barrier, max_over_ranks)
    except Exception as e:                                   # noqa: BLE001
        beam = {"error": f"{type(e).__name__}: {e}"}
    trainer.close()                                        # graphs with captured NCCL work must be gone before the process group is
    if rank != 0:
        return
    cpu = None
    if world == 1 and not skip_side:
        threads = os.cpu_count() or 1
        v, cms, kind, sample = cpu_reference_rate(CPU_SAMPLE_B, 3, 1, threads)
        cpu = {"value": v, "unit": UNIT, "cores": threads, "kind": kind, "sample": sample}
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
            "data": "synthetic",
            "config": {"workload": WORKLOAD, "global_batch": world * B, "parallelism": f"dp{world}",
                       "l2": "inputs larger than L2 (a+Wa = 257 MB per step > 126 MB), two alternating batches",
                       "optimizer": "fused Adam inside the timed step", "dropout": "train mode (p=0.5)", "loss": loss_val,
                       "launch": graph_note or ("one CUDA graph replay per step (mask, forward, loss, backward, all-reduce, Adam captured)"
                                                if trainer.use_cuda_graph else "eager kernel launches"),
                       "grad_exchange": None if world == 1 else ((f"bf16 all-reduce (NCCL over NVLink), {trainer.buckets} buckets, the early one(s) under the backward tail" if trainer.overlap_allreduce
                                                                   else "bf16 all-reduce (NCCL over NVLink), one exchange after the backward pass")
                                                                  if gx == torch.bfloat16 else f"fp32 all-reduce (NCCL over NVLink), {trainer.buckets} buckets"),
                       "repeats": f"median of {args.repeats} timed regions of {args.steps} steps", "repeats_ms_per_step": reps,
                       "host_numa": numa},
            "clocks": clocks,
            "e2e": {"value": e2e_val, "unit": UNIT, "ms_per_step": e2e_ms, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "mode": "features uploaded from pinned host memory every step", "repeats_ms_per_step": e2e_reps,
                    "h2d_alone_ms": h2d_ms, "h2d_alone_gbs": host[0][0].numel() * 2 / h2d_ms / 1e6},
            "e2e_cached": {"value": e2ec_val, "unit": UNIT, "ms_per_step": e2ec_ms, "h2d_bytes_per_step": h2d_cached,
                           "d2h_bytes_per_step": d2h, "repeats_ms_per_step": e2ec_reps,
                           "mode": "features gathered from the HBM-resident FeatureCache (frozen encoder, encoder.py:29-31): the step "
                                   "uploads batch indices + captions from pinned host memory"},
            "gpu_launches": None if launches is None else launches * args.steps,
            "gpu_launches_per_step": launches,
            "roofline": roofline, "cpu_baseline": cpu, "beam": beam}
    print(json.dumps(line), file=_JSON_OUT or sys.stdout, flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--repeats", type=int, default=3, help="timed regions of --steps steps; the median is reported")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    args.warmup = max(args.warmup, 3)
    args.repeats = max(args.repeats, 1)
    # libraries (NCCL's version banner) write to fd 1: keep the real stdout for the ONE JSON line, send the rest to stderr
    global _JSON_OUT
    sys.stdout.flush()
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    try:
        run_ours(args, rank, world, local_rank)
    except BaseException:
        # a rank that dies while its peers sit in a collective would hang the job (and destroy_process_group() hangs while CUDA
        # graphs with captured NCCL work are alive in the failed frame): report and leave at once
        import traceback
        traceback.print_exc()
        sys.stderr.flush()
        os._exit(1)
    if world > 1:
        import torch.distributed as dist
        torch.cuda.synchronize()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
