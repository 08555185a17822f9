"""Drop-in `Decoder` for the reference's decoder.py (same constructor, parameter names, state_dict
keys and return types), backed by the sm_100a kernels in libsat_b200.so through a C ABI.

Reference surface mirrored here (paths relative to the upstream repo):
  * Decoder.__init__            decoder.py:10-67
  * Decoder.forward             decoder.py:69-135   -> (preds (B,T,V) fp32, alphas (B,T,L) fp32)
  * Decoder.get_init_lstm_state decoder.py:137-147
  * Decoder.caption             decoder.py:160-269  -> (list[int], list[list[float]]) or ([0], Tensor)

Extensions (keyword-only, defaulting to reference behaviour): `compute_dtype` (torch.float32 or
torch.bfloat16), `gate` (False = "no gating scalar"), `forward_loss` (fused train.py loss),
`caption_batch` (batched beam search).  There is no CPU fallback: inputs must live on a CUDA device.
"""
from __future__ import annotations

import ctypes as C

import torch
import torch.nn as nn

from . import _lib
from .attention import Attention


def _fill_params(struct, module: "Decoder", grads: bool = False, tensors=None):
    """Point the C struct's parameter fields at the module's fp32 parameter storage."""
    keep = []
    sd = dict(module.named_parameters())
    for field in _lib.PARAM_FIELDS:
        key = _lib.PARAM_KEYS[field]
        t = tensors.get(key) if tensors is not None else sd.get(key)
        if t is None:
            setattr(struct, field, None)
            continue
        if not grads:
            if t.dtype != torch.float32 or not t.is_contiguous():
                raise RuntimeError(f"sat_b200: parameter {key} must be contiguous float32")
        keep.append(t)
        setattr(struct, field, t.data_ptr())
    return keep


_PINNED = {}


def _to_host(tensors):
    """Device tensors -> host tensors with ONE synchronisation: non-blocking copies into cached pinned buffers (six separate
    `.cpu()` calls of pageable copies cost ~1 ms per beam-search call, r2 beam timeline).  The returned tensors are views of the
    cache: consume them before the next call (results_from_raw converts them to lists / clones at once)."""
    out, dev = {}, None
    for k, t in tensors.items():
        if not t.is_cuda:
            out[k] = t
            continue
        dev = t.device
        key = (k, tuple(t.shape), t.dtype)
        buf = _PINNED.get(key)
        if buf is None:
            if len(_PINNED) > 64:
                _PINNED.clear()
            buf = _PINNED[key] = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
        buf.copy_(t, non_blocking=True)
        out[k] = buf
    if dev is not None:
        torch.cuda.current_stream(dev).synchronize()
    return out


class _DecoderFn(torch.autograd.Function):
    """autograd bridge: forward = sat_decoder_forward, backward = sat_decoder_backward."""

    @staticmethod
    def forward(ctx, module, img_features, captions, dropout_mask, fused, alpha_c, need_grad, *params):
        L = _lib.lib()
        dev = img_features.device
        B, Lloc, D = img_features.shape
        T = captions.shape[1] - 1
        H, V = module.embedding_size, module.vocabulary_size
        a = _lib.DecoderArgs()
        a.struct_bytes = C.sizeof(_lib.DecoderArgs)
        a.dtype = _lib.dtype_code(module.compute_dtype)
        a.B, a.L, a.D, a.H, a.V, a.T = B, Lloc, D, H, V, T
        a.tf, a.ado, a.attention = int(module.use_tf), int(module.use_advanced_deep_output), int(module.use_attention)
        a.gate = int(module.use_gate)
        a.start_token = int(module.start_token)
        a.need_grad = int(need_grad)
        # train.py:132 runs the encoder outside no_grad: the gradient reaches img_features when it requires grad
        need_dfeat = bool(need_grad and ctx.needs_input_grad[1])
        a.need_dfeat = int(need_dfeat)
        a.fused_loss = int(fused)
        a.alpha_c = float(alpha_c)
        a.a_dtype = _lib.dtype_code(img_features.dtype)
        a.img_features = img_features.data_ptr()
        a.a_stride_b, a.a_stride_l, a.a_stride_d = img_features.stride()
        a.captions = captions.data_ptr()
        a.cap_stride = captions.stride(0)
        a.dropout_mask = None if dropout_mask is None else dropout_mask.data_ptr()
        keep = _fill_params(a, module)
        preds = torch.empty(B, T, V, device=dev, dtype=torch.float32)          # decoder.py:92
        alphas = torch.empty(B, T, Lloc, device=dev, dtype=torch.float32)      # decoder.py:93
        loss = torch.zeros(4, device=dev, dtype=torch.float32)          # [ce, alpha regulariser, total, status bits]
        a.preds, a.alphas, a.loss = preds.data_ptr(), alphas.data_ptr(), loss.data_ptr()
        nbytes = C.c_int64(0)
        with torch.cuda.device(dev):
            _lib.check(L.sat_decoder_workspace_bytes(C.byref(a), C.byref(nbytes)), "sat_decoder_workspace_bytes")
            ws = torch.empty(nbytes.value + 256, device=dev, dtype=torch.uint8)
            off = (-ws.data_ptr()) % 256
            a.workspace = ws.data_ptr() + off
            a.workspace_bytes = nbytes.value
            a.stream = torch.cuda.current_stream(dev).cuda_stream
            _lib.check(L.sat_decoder_forward(C.byref(a)), "sat_decoder_forward")
        ctx.need_dfeat = need_dfeat
        ctx.args = a
        ctx.keep = (keep, ws, img_features, captions, dropout_mask, preds, alphas, loss)
        ctx.module = module
        ctx.param_list = params
        ctx.fused = fused
        ctx.set_materialize_grads(False)
        if fused:
            ctx.mark_non_differentiable(preds, alphas)    # the regulariser on alphas is already inside the fused loss
        if fused:
            return loss[2], preds, alphas
        return preds, alphas

    @staticmethod
    def backward(ctx, *gouts):
        L = _lib.lib()
        module = ctx.module
        a = ctx.args
        dev = ctx.keep[2].device
        g = _lib.DecoderGrads()
        g.struct_bytes = C.sizeof(_lib.DecoderGrads)
        hold = []
        if ctx.fused:
            dloss = gouts[0]
            if dloss is None:
                dloss = torch.zeros(1, device=dev)
            dloss = dloss.to(torch.float32).contiguous().reshape(1)
            hold.append(dloss)
            g.dloss = dloss.data_ptr()
            g.dpreds = None
        else:
            dpreds, dalphas = gouts
            if dpreds is None:
                dpreds = torch.zeros_like(ctx.keep[5])
            dpreds = dpreds.to(torch.float32).contiguous()
            hold.append(dpreds)
            g.dpreds = dpreds.data_ptr()
            if dalphas is not None and module.use_attention:
                dalphas = dalphas.to(torch.float32).contiguous()
                hold.append(dalphas)
                g.dalphas = dalphas.data_ptr()
        named = dict(module.named_parameters())
        grads = {}
        unused = module.unused_parameter_keys()
        for field in _lib.PARAM_FIELDS:
            key = _lib.PARAM_KEYS[field]
            p = named.get(key)
            if p is None or not p.requires_grad or key in unused:
                setattr(g, field, None)
                continue
            gt = torch.empty_like(p, dtype=torch.float32)
            grads[key] = gt
            setattr(g, field, gt.data_ptr())
        dfeat = None
        if ctx.need_dfeat:
            feats = ctx.keep[2]
            dfeat = torch.empty(feats.shape, device=dev, dtype=torch.float32)
            g.img_features = dfeat.data_ptr()
        with torch.cuda.device(dev):
            a.stream = torch.cuda.current_stream(dev).cuda_stream
            _lib.check(L.sat_decoder_backward(C.byref(a), C.byref(g)), "sat_decoder_backward")
        if dfeat is not None and dfeat.dtype != ctx.keep[2].dtype:
            dfeat = dfeat.to(ctx.keep[2].dtype)
        out = []
        for name, p in module.named_parameters():
            out.append(grads.get(name))       # None reproduces the reference's grad=None pattern (Q6, Q7)
        return (None, dfeat, None, None, None, None, None, *out)


class Decoder(nn.Module):
    def __init__(self, vocabulary_size, encoder_dim, tf=False, ado=False, bert=False, attention=False, *,
                 compute_dtype=torch.float32, gate=True):
        super().__init__()
        self.use_tf = tf
        self.use_advanced_deep_output = ado
        self.use_bert = bert
        self.use_attention = attention
        self.use_gate = gate
        self.compute_dtype = compute_dtype
        self.encoder_dim = encoder_dim
        self.validate_tokens = True

        if bert:  # decoder.py:21-36
            from transformers import BertModel, BertTokenizer
            bert_model = BertModel.from_pretrained('bert-base-uncased')
            self.tokenizer = BertTokenizer.from_pretrained('bert-base-uncased')
            self.vocabulary_size = bert_model.config.vocab_size
            self.embedding_size = bert_model.config.hidden_size
            self.embedding = bert_model.get_input_embeddings()
            for param in self.embedding.parameters():
                param.requires_grad = False
            del bert_model
            self.start_token = self.tokenizer.cls_token_id
        else:     # decoder.py:38-40
            self.vocabulary_size = vocabulary_size
            self.embedding_size = 512
            self.embedding = nn.Embedding(self.vocabulary_size, self.embedding_size)
            self.start_token = 0
        H = self.embedding_size
        self.init_h = nn.Linear(encoder_dim, H)            # decoder.py:43-44
        self.init_c = nn.Linear(encoder_dim, H)
        self.tanh = nn.Tanh()
        self.f_beta = nn.Linear(H, encoder_dim)            # decoder.py:48
        self.sigmoid = nn.Sigmoid()
        self.attention = Attention(encoder_dim, H, compute_dtype=compute_dtype)   # decoder.py:52
        self.lstm = nn.LSTMCell(H + encoder_dim, H)        # decoder.py:53
        if ado:                                            # decoder.py:56-63
            self.f_h = nn.Linear(H, H)
            self.f_z = nn.Linear(encoder_dim, H)
            self.f_out = nn.Linear(H, self.vocabulary_size)
            self.relu = nn.ReLU()
            self.dropout = nn.Dropout()
        self.deep_output = nn.Linear(H, self.vocabulary_size)   # decoder.py:66 (always constructed, Q7)
        self.dropout = nn.Dropout()

    # ------------------------------------------------------------------------------------------
    def unused_parameter_keys(self):
        """Parameters the reference leaves with grad=None."""
        unused = set()
        if self.use_advanced_deep_output:
            unused |= {"deep_output.weight", "deep_output.bias"}
        if not self.use_attention:
            unused |= {"attention.U.weight", "attention.U.bias", "attention.W.weight", "attention.W.bias",
                       "attention.v.weight", "attention.v.bias", "f_beta.weight", "f_beta.bias"}
        elif not self.use_gate:
            unused |= {"f_beta.weight", "f_beta.bias"}
        return unused

    def _check_inputs(self, img_features, captions=None, check_tokens=None):
        _lib.require_cuda(img_features, "img_features")
        if img_features.dim() != 3 or img_features.shape[2] != self.encoder_dim:
            raise RuntimeError(f"img_features must be (B, L, {self.encoder_dim}); got {tuple(img_features.shape)}")
        if img_features.dtype not in (torch.float32, torch.bfloat16):
            raise TypeError("img_features must be float32 or bfloat16")
        if captions is not None:
            if captions.dtype != torch.int64:
                captions = captions.long()
            if captions.device != img_features.device:
                captions = captions.to(img_features.device)
            if captions.stride(1) != 1:
                captions = captions.contiguous()
            if self.validate_tokens and check_tokens is not None and captions.shape[1] > 1:
                # nn.Embedding (decoder.py:87) / CrossEntropyLoss (train.py:161) raise IndexError on an id outside [0, V); the
                # kernels only flag it (loss[3]) and read a zero row.  Columns the reference touches: [0, T) as step inputs
                # under teacher forcing (free-running decoding uses `captions` for its length only, decoder.py:77), [1, T) as
                # targets of the fused loss.  One tiny reduction + host sync per call; `decoder.validate_tokens = False`
                # skips it (FusedTrainStep checks the device flag lazily instead).
                used = captions[:, :-1] if self.use_tf else (captions[:, 1:-1] if check_tokens == "loss" else None)
                if used is not None and used.numel():
                    lo, hi = torch.aminmax(used)
                    if int(lo) < 0 or int(hi) >= self.vocabulary_size:
                        raise IndexError(f"caption token out of range [0, {self.vocabulary_size}): min {int(lo)}, max {int(hi)}")
        return captions

    def _dropout_mask(self, B, T, device, dropout_masks):
        if dropout_masks is not None:
            return dropout_masks.to(device=device, dtype=torch.float32).contiguous()
        if not self.training:
            return None
        p = self.dropout.p
        if p <= 0.0:
            return None
        keep = torch.rand(B, T, self.embedding_size, device=device) >= p        # decoder.py:121-125 (Q9)
        return keep.to(torch.float32) / (1.0 - p)

    def forward(self, img_features, captions, *, dropout_masks=None):
        captions = self._check_inputs(img_features, captions, "forward")
        B, T = img_features.shape[0], captions.shape[1] - 1
        if B == 0 or T <= 0:
            # decoder.py:92-93 allocate zero tensors and the `for t in range(max_timespan)` loop (decoder.py:96) never runs: captions
            # holding only <start>, or an empty batch, give empty preds / alphas that no parameter reaches
            T = max(T, 0)
            return (torch.zeros(B, T, self.vocabulary_size, device=img_features.device),
                    torch.zeros(B, T, img_features.shape[1], device=img_features.device))
        mask = self._dropout_mask(B, T, img_features.device, dropout_masks)
        params = tuple(p for _, p in self.named_parameters())
        need_grad = torch.is_grad_enabled() and (img_features.requires_grad or any(p.requires_grad for p in params))
        return _DecoderFn.apply(self, img_features, captions, mask, False, 0.0, need_grad, *params)

    def forward_loss(self, img_features, captions, alpha_c=1.0, *, dropout_masks=None):
        """Extension: forward + the train.py:150-162 loss fused on device.  Returns (loss, preds, alphas)."""
        captions = self._check_inputs(img_features, captions, "loss")
        B, T = img_features.shape[0], captions.shape[1] - 1
        mask = self._dropout_mask(B, T, img_features.device, dropout_masks)
        params = tuple(p for _, p in self.named_parameters())
        need_grad = torch.is_grad_enabled() and (img_features.requires_grad or any(p.requires_grad for p in params))
        return _DecoderFn.apply(self, img_features, captions, mask, True, float(alpha_c), need_grad, *params)

    # ------------------------------------------------------------------------------------------
    def get_init_lstm_state(self, img_features):
        """decoder.py:137-147 through sat_gemm (tanh epilogue).  Inference helper (no autograd)."""
        self._check_inputs(img_features)
        L = _lib.lib()
        dev = img_features.device
        dt = self.compute_dtype
        code = _lib.dtype_code(dt)
        with torch.no_grad(), torch.cuda.device(dev):
            avg = img_features.to(torch.float32).mean(dim=1).to(dt).contiguous()
            w = torch.cat((self.init_h.weight, self.init_c.weight), 0).to(dt).contiguous()
            b = torch.cat((self.init_h.bias, self.init_c.bias), 0).float().contiguous()
            B, D, H = avg.shape[0], self.encoder_dim, self.embedding_size
            out = torch.empty(B, 2 * H, device=dev, dtype=torch.float32)
            _lib.check(L.sat_gemm(code, B, 2 * H, D, avg.data_ptr(), D, 1, w.data_ptr(), D, 1, b.data_ptr(), 1,
                                  out.data_ptr(), 2 * H, None, 0, 0, torch.cuda.current_stream(dev).cuda_stream), "sat_gemm")
        return out[:, :H].contiguous(), out[:, H:].contiguous()

    # ------------------------------------------------------------------------------------------
    def caption_batch_raw(self, img_features, beam_size, max_steps=50):
        """Batched beam search, device-side outputs only (no host sync): dict of lens / status / live (NI,) int32, tokens
        (NI, max_steps+2) int64, alphas (NI, max_steps+2, L) fp32, last_alpha (NI, beam, L) fp32."""
        self._check_inputs(img_features)
        L = _lib.lib()
        dev = img_features.device
        NI, Lloc, D = img_features.shape
        H, V = self.embedding_size, self.vocabulary_size
        a = _lib.BeamArgs()
        a.struct_bytes = C.sizeof(_lib.BeamArgs)
        a.dtype = _lib.dtype_code(self.compute_dtype)
        a.NI, a.L, a.D, a.H, a.V = NI, Lloc, D, H, V
        a.beam, a.max_steps = int(beam_size), int(max_steps)
        a.ado, a.attention, a.gate, a.bert = (int(self.use_advanced_deep_output), int(self.use_attention),
                                               int(self.use_gate), int(self.use_bert))
        a.start_token = int(self.start_token)
        a.a_dtype = _lib.dtype_code(img_features.dtype)
        a.img_features = img_features.data_ptr()
        a.a_stride_b, a.a_stride_l, a.a_stride_d = img_features.stride()
        keep = _fill_params(a, self)
        maxlen = max_steps + 2
        with torch.cuda.device(dev):
            out_len = torch.zeros(NI, device=dev, dtype=torch.int32)
            out_status = torch.zeros(NI, device=dev, dtype=torch.int32)
            out_live = torch.zeros(NI, device=dev, dtype=torch.int32)
            out_tokens = torch.zeros(NI, maxlen, device=dev, dtype=torch.int64)
            out_alphas = torch.zeros(NI, maxlen, Lloc, device=dev, dtype=torch.float32)
            out_last = torch.zeros(NI, beam_size, Lloc, device=dev, dtype=torch.float32)
            a.out_len, a.out_status, a.out_tokens = out_len.data_ptr(), out_status.data_ptr(), out_tokens.data_ptr()
            a.out_alphas, a.out_last_alpha, a.out_live = out_alphas.data_ptr(), out_last.data_ptr(), out_live.data_ptr()
            nbytes = C.c_int64(0)
            _lib.check(L.sat_beam_workspace_bytes(C.byref(a), C.byref(nbytes)), "sat_beam_workspace_bytes")
            ws = torch.empty(nbytes.value + 256, device=dev, dtype=torch.uint8)
            a.workspace = ws.data_ptr() + ((-ws.data_ptr()) % 256)
            a.workspace_bytes = nbytes.value
            a.stream = torch.cuda.current_stream(dev).cuda_stream
            with torch.no_grad():
                _lib.check(L.sat_beam_search(C.byref(a)), "sat_beam_search")
        del keep
        return {"lens": out_len, "status": out_status, "live": out_live, "tokens": out_tokens, "alphas": out_alphas,
                "last_alpha": out_last}

    @staticmethod
    def results_from_raw(raw):
        """Padded device outputs -> the reference's per-image return values: (list[int], list[list[float]]) for a completed
        caption, ([0], Tensor(live, L)) for the "No completed sentences found" fallback (decoder.py:256-258, Q18)."""
        host = _to_host({k: raw[k] for k in ("lens", "status", "live", "tokens", "alphas", "last_alpha") if k in raw})
        lens, status, live = host["lens"].tolist(), host["status"].tolist(), host["live"].tolist()
        toks, alph = host["tokens"], host.get("alphas")
        last = host.get("last_alpha") if any(status) else None     # fallback rows (the reference builds these on the CPU, Q19)
        results = []
        for i in range(len(lens)):
            if status[i] != 0:
                results.append(([0], last[i, :live[i]].clone() if last is not None else None))
            else:
                n = lens[i]
                results.append((toks[i, :n].tolist(), alph[i, :n].tolist() if alph is not None else None))
        return results

    def caption_batch(self, img_features, beam_size, max_steps=50):
        """Extension: beam search for NI images at once; img_features is (NI, L, D), one row per image.
        Returns a list of (sentence, alpha) pairs with the reference's per-image return convention."""
        return self.results_from_raw(self.caption_batch_raw(img_features, beam_size, max_steps))

    def caption(self, img_features, beam_size):
        """decoder.py:160-269.  `img_features` is one image expanded to (beam, L, D) as in
        generate_caption.py:89; row 0 is the image."""
        sentence, alpha = self.caption_batch(img_features[:1], beam_size)[0]
        if sentence == [0] and torch.is_tensor(alpha):
            print('No completed sentences found')                # decoder.py:257
        return sentence, alpha
