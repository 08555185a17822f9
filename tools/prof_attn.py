"""Tiny driver for ncu: a few attention-step fwd/bwd launches at the c2 shape through the C ABI."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sat_b200 import _lib  # noqa: E402

B, L, D, H = 256, 196, 2048, 512
dt = torch.bfloat16
Lb = _lib.lib()
g = torch.Generator(device="cuda").manual_seed(0)
a = torch.relu(torch.randn(B, L, D, device="cuda", generator=g)).to(dt)
wa = torch.randn(B, L, H, device="cuda", generator=g).to(dt)
uf = torch.randn(B, H + D, device="cuda", generator=g)
v = torch.randn(H, device="cuda", generator=g) * 0.05
bv = torch.zeros(1, device="cuda")
alpha = torch.empty(B, L, device="cuda")
ctx = torch.empty(B, D, device="cuda", dtype=dt)
z = torch.empty(B, D, device="cuda", dtype=dt)
dz = torch.randn(B, D, device="cuda", generator=g)
duf = torch.empty(B, H + D, device="cuda", dtype=dt)
dwa = torch.zeros(B, L, H, device="cuda")
dv = torch.zeros(B, H, device="cuda")
dbv = torch.zeros(B, device="cuda")
st = torch.cuda.current_stream().cuda_stream
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4
e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
for it in range(n):
    if it == n - 1:
        e[0].record()
    _lib.check(Lb.sat_attn_step_fwd(1, B, 1, L, D, H, a.data_ptr(), wa.data_ptr(), uf.data_ptr(), H + D, v.data_ptr(), bv.data_ptr(),
                                    1, alpha.data_ptr(), L, ctx.data_ptr(), D, z.data_ptr(), D, st))
    if it == n - 1:
        e[1].record()
    _lib.check(Lb.sat_attn_step_bwd(1, B, L, D, H, a.data_ptr(), wa.data_ptr(), uf.data_ptr(), H + D, v.data_ptr(), alpha.data_ptr(), L,
                                    ctx.data_ptr(), D, dz.data_ptr(), D, None, 0, None, 0, 1, duf.data_ptr(), H + D, dwa.data_ptr(),
                                    dv.data_ptr(), dbv.data_ptr(), None, 0, st))
    if it == n - 1:
        e[2].record()
torch.cuda.synchronize()
print("fwd ms %.4f  bwd ms %.4f" % (e[0].elapsed_time(e[1]), e[1].elapsed_time(e[2])))
