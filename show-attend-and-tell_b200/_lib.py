"""ctypes binding of libsat_b200.so (C ABI declared in include/sat_b200.h).

The library is the product path: there is no Python/PyTorch fallback.  `lib()` raises if the
shared object is missing (run `python __graft_entry__.py` or `python show-attend-and-tell_b200/build.py`).
"""
from __future__ import annotations

import ctypes as C
import os

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "lib", "libsat_b200.so")

DTYPE_F32, DTYPE_BF16 = 0, 1
_vp, _i32, _i64, _f32 = C.c_void_p, C.c_int32, C.c_int64, C.c_float

PARAM_FIELDS = [
    "embedding", "init_h_w", "init_h_b", "init_c_w", "init_c_b", "f_beta_w", "f_beta_b",
    "att_U_w", "att_U_b", "att_W_w", "att_W_b", "att_v_w", "att_v_b",
    "lstm_w_ih", "lstm_w_hh", "lstm_b_ih", "lstm_b_hh", "deep_w", "deep_b",
    "f_h_w", "f_h_b", "f_z_w", "f_z_b", "f_out_w", "f_out_b",
]
# C field name -> reference state_dict key
PARAM_KEYS = {
    "embedding": "embedding.weight",
    "init_h_w": "init_h.weight", "init_h_b": "init_h.bias", "init_c_w": "init_c.weight", "init_c_b": "init_c.bias",
    "f_beta_w": "f_beta.weight", "f_beta_b": "f_beta.bias",
    "att_U_w": "attention.U.weight", "att_U_b": "attention.U.bias",
    "att_W_w": "attention.W.weight", "att_W_b": "attention.W.bias",
    "att_v_w": "attention.v.weight", "att_v_b": "attention.v.bias",
    "lstm_w_ih": "lstm.weight_ih", "lstm_w_hh": "lstm.weight_hh", "lstm_b_ih": "lstm.bias_ih", "lstm_b_hh": "lstm.bias_hh",
    "deep_w": "deep_output.weight", "deep_b": "deep_output.bias",
    "f_h_w": "f_h.weight", "f_h_b": "f_h.bias", "f_z_w": "f_z.weight", "f_z_b": "f_z.bias",
    "f_out_w": "f_out.weight", "f_out_b": "f_out.bias",
}


class DecoderArgs(C.Structure):
    _fields_ = (
        [("struct_bytes", _i32), ("dtype", _i32)]
        + [(n, _i32) for n in ("B", "L", "D", "H", "V", "T", "tf", "ado", "attention", "gate", "start_token",
                               "need_grad", "fused_loss", "a_dtype")]
        + [("alpha_c", _f32), ("need_dfeat", _i32)]
        + [("img_features", _vp), ("a_stride_b", _i64), ("a_stride_l", _i64), ("a_stride_d", _i64),
           ("captions", _vp), ("cap_stride", _i64), ("dropout_mask", _vp)]
        + [(n, _vp) for n in PARAM_FIELDS]
        + [("preds", _vp), ("alphas", _vp), ("loss", _vp), ("workspace", _vp), ("workspace_bytes", _i64), ("stream", _vp)]
    )


class DecoderGrads(C.Structure):
    _fields_ = (
        [("struct_bytes", _i32), ("reserved", _i32), ("dpreds", _vp), ("dalphas", _vp), ("dloss", _vp)]
        + [(n, _vp) for n in PARAM_FIELDS]
        + [("img_features", _vp)]
    )


class BeamArgs(C.Structure):
    _fields_ = (
        [("struct_bytes", _i32), ("dtype", _i32)]
        + [(n, _i32) for n in ("NI", "L", "D", "H", "V", "beam", "max_steps", "ado", "attention", "gate", "bert",
                               "start_token", "a_dtype")]
        + [("img_features", _vp), ("a_stride_b", _i64), ("a_stride_l", _i64), ("a_stride_d", _i64)]
        + [(n, _vp) for n in PARAM_FIELDS]
        + [("out_len", _vp), ("out_status", _vp), ("out_tokens", _vp), ("out_alphas", _vp), ("out_last_alpha", _vp),
           ("out_live", _vp), ("workspace", _vp), ("workspace_bytes", _i64), ("stream", _vp)]
    )


_LIB = None

EXPORTS = ["sat_last_error", "sat_version", "sat_set_force_simt", "sat_set_fp32_gemm_mode", "sat_set_gemm_wide_tile", "sat_set_attn_generic", "sat_set_attn_partition", "sat_set_pdl", "sat_set_gemm_static_weights", "sat_set_tail_overlap", "sat_set_dpre_recompute", "sat_set_fused_cell_backward", "sat_gemm_phase_stamps", "sat_set_decoder_split", "sat_set_decoder_split_sms", "sat_set_gemm_splitk", "sat_set_attn_l2_policy", "sat_set_attn_prefetch", "sat_gemm", "sat_attn_step_fwd", "sat_attn_step_fwd_static", "sat_attn_step_bwd", "sat_attn_scratch_bytes",
           "sat_decoder_workspace_bytes", "sat_decoder_logits", "sat_wait_early_grads", "sat_wait_side_grads", "sat_logit_rank", "sat_decoder_forward", "sat_decoder_backward", "sat_beam_workspace_bytes",
           "sat_beam_search", "sat_adam_step", "sat_adam_step_dev", "sat_adam_step_dev_bf16grad", "sat_cast_f32_to_bf16", "sat_adam_step_dev_part", "sat_count_tokens", "sat_argmax_rows", "sat_gather_items",
           "sat_profile", "sat_profile_read"]


def lib():
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"sat_b200: {LIB_PATH} is missing. Build it with `python show-attend-and-tell_b200/build.py` "
                "(nvcc, sm_100a). There is no CPU / PyTorch fallback for this path.")
        L = C.CDLL(LIB_PATH)
        L.sat_last_error.restype = C.c_char_p
        L.sat_version.restype = C.c_int
        L.sat_gemm.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, _vp, _i64, C.c_int, _vp, _i64, C.c_int, _vp, C.c_int,
                               _vp, _i64, _vp, _i64, C.c_int, _vp]
        L.sat_attn_step_fwd.argtypes = [C.c_int] * 6 + [_vp, _vp, _vp, _i64, _vp, _vp, C.c_int, _vp, _i64, _vp, _i64,
                                                        _vp, _i64, _vp, _i64, _vp, _i64, _vp]
        L.sat_attn_step_fwd_static.argtypes = [C.c_int] * 6 + [_vp, _vp, _vp, _i64, _vp, _vp, C.c_int, _vp, _i64, _vp, _i64,
                                                               _vp, _i64, _vp, _i64, _vp]
        L.sat_set_attn_l2_policy.argtypes = [C.c_int, C.c_int, _i64]
        L.sat_attn_scratch_bytes.argtypes = [C.c_int] * 5 + [C.POINTER(_i64)]
        L.sat_attn_step_bwd.argtypes = ([C.c_int] * 5 + [_vp, _vp, _vp, _i64, _vp, _vp, _i64, _vp, _i64, _vp, _i64, _vp,
                                                         _i64, _vp, _i64, C.c_int, _vp, _i64, _vp, _vp, _vp, _vp, C.c_int, _vp, _vp, _i64, _vp, _i64, _vp])
        L.sat_decoder_workspace_bytes.argtypes = [C.POINTER(DecoderArgs), C.POINTER(_i64)]
        L.sat_decoder_logits.argtypes = [C.POINTER(DecoderArgs), C.POINTER(_vp), C.POINTER(_i64)]
        L.sat_wait_early_grads.argtypes = [_vp]
        L.sat_wait_side_grads.argtypes = [_vp]
        L.sat_adam_step_dev_part.argtypes = [_vp, _vp, C.c_int, _vp, _vp, _i64, _vp, _vp, C.c_int, _vp]
        L.sat_logit_rank.argtypes = [_vp, _i64, _i64, C.c_int, _vp, _vp, _vp]
        L.sat_decoder_forward.argtypes = [C.POINTER(DecoderArgs)]
        L.sat_decoder_backward.argtypes = [C.POINTER(DecoderArgs), C.POINTER(DecoderGrads)]
        L.sat_beam_workspace_bytes.argtypes = [C.POINTER(BeamArgs), C.POINTER(_i64)]
        L.sat_beam_search.argtypes = [C.POINTER(BeamArgs)]
        L.sat_adam_step.argtypes = [_vp, _vp, _vp, _vp, _i64, _f32, _f32, _f32, _f32, _f32, C.c_int, _f32, _vp]
        L.sat_adam_step_dev.argtypes = [_vp, _vp, _vp, _vp, _i64, _vp, _vp, _vp]
        L.sat_adam_step_dev_bf16grad.argtypes = [_vp, _vp, _vp, _vp, _i64, _vp, _vp, _vp]
        L.sat_cast_f32_to_bf16.argtypes = [_vp, _vp, _i64, _vp]
        L.sat_count_tokens.argtypes = [_vp, _i64, _i64, _i64, _i64, _i64, _i64, _vp, _vp]
        L.sat_argmax_rows.argtypes = [_vp, _i64, _i64, C.c_int, _vp, _vp]
        L.sat_gather_items.argtypes = [_vp, _i64, _i64, _vp, _i64, _vp, _vp, _vp]
        _LIB = L
    return _LIB


def attn_scratch(NI, NB, L, D, H, device):
    """Scratch tensor for the streaming attention kernels (no initialisation needed), or None when the shape runs on
    the generic kernels."""
    n = _i64(0)
    check(lib().sat_attn_scratch_bytes(NI, NB, L, D, H, C.byref(n)), "sat_attn_scratch_bytes")
    if n.value == 0:
        return None
    return torch.empty(n.value, device=device, dtype=torch.uint8)


def check(rc: int, what: str = "sat_b200"):
    if rc != 0:
        msg = lib().sat_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"{what} failed (code {rc}): {msg}")


def dtype_code(dt: torch.dtype) -> int:
    if dt == torch.float32:
        return DTYPE_F32
    if dt == torch.bfloat16:
        return DTYPE_BF16
    raise TypeError(f"sat_b200 supports float32 and bfloat16, got {dt}")


def ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def stream_ptr(device):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def require_cuda(t: torch.Tensor, name: str):
    if not t.is_cuda:
        raise RuntimeError(f"sat_b200: {name} must be a CUDA tensor (the hot path has no CPU fallback); got device {t.device}")
