"""Drop-in `Attention` for the reference's attention.py (attention.py:5-21).

Same constructor and parameter names as the reference (note SURVEY Q2: `U` acts on the hidden state,
`W` on the image features).  forward(img_features, hidden_state) -> (context (B,D), alpha (B,L)) runs
  K1  Wa = W a + b_W        tensor-core / fp32 GEMM   (sat_gemm)
      Uh = U h + b_U        GEMM                       (sat_gemm)
  K2  fused score/softmax/context kernel               (sat_attn_step_fwd)
and its autograd backward uses sat_attn_step_bwd plus GEMMs.  Inside `Decoder` these kernels are
driven by the C sequence driver instead (Wa hoisted out of the time loop, Q3).
"""
from __future__ import annotations

import torch
import torch.nn as nn

from . import _lib


def _gemm(dt, M, N, K, A, lda, a_k, B, ldb, b_k, bias=None, act=0, Cf=None, ldcf=0, Ct=None, ldct=0, stream=None):
    L = _lib.lib()
    _lib.check(L.sat_gemm(_lib.dtype_code(dt), M, N, K, A.data_ptr(), lda, int(a_k), B.data_ptr(), ldb, int(b_k),
                          None if bias is None else bias.data_ptr(), act,
                          None if Cf is None else Cf.data_ptr(), ldcf, None if Ct is None else Ct.data_ptr(), ldct, 0,
                          stream), "sat_gemm")


class _AttentionFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, module, a_in, h_in, Uw, Ub, Ww, Wb, vw, vb):
        L = _lib.lib()
        dt = module.compute_dtype
        dev = a_in.device
        st = torch.cuda.current_stream(dev).cuda_stream
        B, Lloc, D = a_in.shape
        H = Uw.shape[0]
        a = a_in.to(dt).contiguous()
        h = h_in.to(dt).contiguous()
        Wc, Uc = Ww.detach().to(dt).contiguous(), Uw.detach().to(dt).contiguous()
        wa = torch.empty(B * Lloc, H, device=dev, dtype=dt)
        _gemm(dt, B * Lloc, H, D, a, D, True, Wc, D, True, bias=Wb.detach(), Ct=wa, ldct=H, stream=st)
        uf = torch.empty(B, H, device=dev, dtype=torch.float32)
        _gemm(dt, B, H, H, h, H, True, Uc, H, True, bias=Ub.detach(), Cf=uf, ldcf=H, stream=st)
        alpha = torch.empty(B, Lloc, device=dev, dtype=torch.float32)
        ctxv = torch.empty(B, D, device=dev, dtype=dt)
        vflat = vw.detach().reshape(-1).contiguous()
        # streaming kernels (scratch + fp32 context for their backward) unless d/d img_features is wanted,
        # which only the generic backward kernel produces
        scratch = None if a_in.requires_grad else _lib.attn_scratch(B, 1, Lloc, D, H, dev)
        ctx32 = torch.empty(B, D, device=dev, dtype=torch.float32) if scratch is not None else None
        _lib.check(L.sat_attn_step_fwd(_lib.dtype_code(dt), B, 1, Lloc, D, H, a.data_ptr(), wa.data_ptr(), uf.data_ptr(), H,
                                       vflat.data_ptr(), vb.detach().data_ptr(), 0, alpha.data_ptr(), Lloc,
                                       ctxv.data_ptr(), D, None, 0, _lib.ptr(ctx32), D, _lib.ptr(scratch),
                                       0 if scratch is None else scratch.numel(), st), "sat_attn_step_fwd")
        ctx.saved = (a, h, Wc, Uc, wa, uf, alpha, ctxv, vflat)
        ctx.stream_state = (scratch, ctx32)
        ctx.module = module
        ctx.a_needs_grad = a_in.requires_grad
        ctx.in_dtypes = (a_in.dtype, h_in.dtype)
        return ctxv.to(a_in.dtype), alpha.to(a_in.dtype) if a_in.dtype != torch.float32 else alpha

    @staticmethod
    def backward(ctx, dctx, dalpha):
        L = _lib.lib()
        module = ctx.module
        dt = module.compute_dtype
        a, h, Wc, Uc, wa, uf, alpha, ctxv, vflat = ctx.saved
        dev = a.device
        st = torch.cuda.current_stream(dev).cuda_stream
        B, Lloc, D = a.shape
        H = Uc.shape[0]
        dz = (torch.zeros(B, D, device=dev) if dctx is None else dctx.to(torch.float32)).contiguous()
        ga = None if dalpha is None else dalpha.to(torch.float32).contiguous()
        duf = torch.zeros(B, H + D, device=dev, dtype=dt)
        scratch, ctx32 = ctx.stream_state
        streaming = scratch is not None
        dwa = torch.empty(B, Lloc, H, device=dev, dtype=dt if streaming else torch.float32)
        dv = torch.empty(B, H, device=dev, dtype=torch.float32)
        dbv = torch.empty(B, device=dev, dtype=torch.float32)
        da = torch.empty(B, Lloc, D, device=dev, dtype=torch.float32) if ctx.a_needs_grad else None
        _lib.check(L.sat_attn_step_bwd(_lib.dtype_code(dt), B, Lloc, D, H, a.data_ptr(), wa.data_ptr(), uf.data_ptr(), H,
                                       vflat.data_ptr(), alpha.data_ptr(), Lloc, ctxv.data_ptr(), D, dz.data_ptr(), D,
                                       None, 0, None if ga is None else ga.data_ptr(), Lloc, 0, duf.data_ptr(), H + D,
                                       None if streaming else dwa.data_ptr(), dv.data_ptr(), dbv.data_ptr(),
                                       None if da is None else da.data_ptr(), 1, dwa.data_ptr() if streaming else None,
                                       _lib.ptr(ctx32), D, _lib.ptr(scratch), 0 if scratch is None else scratch.numel(), st),
                   "sat_attn_step_bwd")
        dUh = duf[:, :H].contiguous()
        dwa_t = dwa.reshape(B * Lloc, H).to(dt)
        dwa = dwa.float()
        gUw = torch.empty(H, H, device=dev, dtype=torch.float32)
        _gemm(dt, H, H, B, dUh, H, False, h, H, False, Cf=gUw, ldcf=H, stream=st)                 # dU = dUh^T h
        gWw = torch.empty(H, D, device=dev, dtype=torch.float32)
        _gemm(dt, H, D, B * Lloc, dwa_t, H, False, a.reshape(B * Lloc, D), D, False, Cf=gWw, ldcf=D, stream=st)   # dW = dWa^T a
        gh = torch.empty(B, H, device=dev, dtype=torch.float32)
        _gemm(dt, B, H, H, dUh, H, True, Uc, H, False, Cf=gh, ldcf=H, stream=st)                  # dh = dUh U
        ga_full = None
        if da is not None:
            t = torch.empty(B * Lloc, D, device=dev, dtype=torch.float32)
            _gemm(dt, B * Lloc, D, H, dwa_t, H, True, Wc, D, False, Cf=t, ldcf=D, stream=st)      # da += dWa W
            ga_full = (da + t.view(B, Lloc, D)).to(ctx.in_dtypes[0])
        return (None, ga_full, gh.to(ctx.in_dtypes[1]), gUw, dUh.float().sum(0), gWw, dwa.sum((0, 1)),
                dv.sum(0, keepdim=True), dbv.sum().reshape(1))


class Attention(nn.Module):
    def __init__(self, encoder_dim, embedding_size, *, compute_dtype=torch.float32):
        super().__init__()
        self.U = nn.Linear(embedding_size, embedding_size)   # attention.py:8  (hidden state)
        self.W = nn.Linear(encoder_dim, embedding_size)      # attention.py:9  (image features)
        self.v = nn.Linear(embedding_size, 1)                # attention.py:10
        self.tanh = nn.Tanh()
        self.softmax = nn.Softmax(1)
        self.compute_dtype = compute_dtype

    def forward(self, img_features, hidden_state):
        _lib.require_cuda(img_features, "img_features")
        _lib.require_cuda(hidden_state, "hidden_state")
        return _AttentionFn.apply(self, img_features, hidden_state, self.U.weight, self.U.bias, self.W.weight,
                                  self.W.bias, self.v.weight, self.v.bias)
