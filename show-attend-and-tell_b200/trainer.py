"""Train step driver for the decoder hot path (what train.py:128-164 does around `decoder(...)`):

    loss = trainer.step(img_features, captions)

= sat_decoder_forward (with the train.py:150-162 loss fused) -> sat_decoder_backward writing gradients straight
into one flat fp32 buffer -> (world > 1) one NCCL all-reduce of that buffer over NVLink -> fused Adam
(sat_adam_step_dev) on the flat parameter buffer.  It bypasses autograd: no tape, no per-tensor optimizer launches.
Data parallelism is one process per GPU, batch-sharded, gradients averaged; parameters the
reference leaves with grad=None (Q6, Q7) are simply not part of the flat buffer.

Every step-dependent scalar (learning rate, Adam step count) lives in device memory, so the whole step -- dropout mask,
forward, loss, backward, all-reduce, Adam -- can be captured ONCE into a CUDA graph per (feature buffer, caption buffer)
pair and replayed with a single launch (`use_cuda_graph=True`): a training loop that reads the loss back every step
(train.py:179) otherwise pays the host cost of ~165 kernel launches at the start of every step.

The gradient wrt img_features is not produced here: train.py:71 hands only decoder.parameters() to Adam, so although
train.py:132 runs the encoder outside no_grad nothing ever consumes that gradient in a training step
(`Decoder.forward` produces it on request for callers that do want it).
"""
from __future__ import annotations

import ctypes as C

import torch
import torch.distributed as dist

from . import _lib


class FusedTrainStep:
    # gradients that are final ~160 us before sat_decoder_backward's last kernel (sat_wait_early_grads): 62 % of the bytes at c2
    EARLY_KEYS = ("init_h.weight", "init_h.bias", "init_c.weight", "init_c.bias", "lstm.weight_ih", "lstm.weight_hh")
    # gradients the MAIN chain of the backward tail writes last (sum over t -> dW_att GEMM -> column sums)
    LAST_KEYS = ("attention.W.weight", "attention.W.bias", "attention.v.weight", "attention.v.bias")

    def __init__(self, decoder, lr=1e-4, betas=(0.9, 0.999), eps=1e-8, alpha_c=1.0, process_group=None,
                 distributed=None, step_size=None, gamma=0.1, use_cuda_graph=False, grad_exchange_dtype=torch.float32):
        self.dec = decoder
        self.base_lr, self.betas, self.eps, self.alpha_c = float(lr), betas, eps, alpha_c
        self.distributed = dist.is_initialized() if distributed is None else distributed
        self.pg = process_group
        self.world = dist.get_world_size(process_group) if self.distributed else 1
        unused = decoder.unused_parameter_keys()
        keys = [n for n, p in decoder.named_parameters() if p.requires_grad and n not in unused]
        # the gradients sat_decoder_backward finishes early go first in the flat buffer: one contiguous all-reduce bucket
        # ... then everything the side chain of the backward tail produces, and last the encoder projection's gradients, which
        # the main chain finishes at the very end (sat_wait_side_grads): three contiguous buckets
        early = [k for k in keys if k in self.EARLY_KEYS]
        last = [k for k in keys if k in self.LAST_KEYS]
        self.train_keys = early + [k for k in keys if k not in self.EARLY_KEYS and k not in self.LAST_KEYS] + last
        named = dict(decoder.named_parameters())
        dev = next(decoder.parameters()).device
        self.device = dev
        offsets, off = {}, 0
        for k in self.train_keys:
            offsets[k] = off
            off = (off + named[k].numel() + 3) & ~3          # keep every tensor 16-byte aligned
        n = off
        self.offsets = offsets
        self.n_early = max([offsets[k] + ((named[k].numel() + 3) & ~3) for k in early], default=0)
        self.n_side = min([offsets[k] for k in last], default=n)          # [n_early, n_side): side-chain gradients; [n_side, n): last
        # Gradient exchange schedule.  overlap_allreduce = False (default): ONE all-reduce of the whole flat buffer after the backward
        # pass.  True: buckets exchanged on a comm stream as sat_decoder_backward finishes them -- buckets = 2: the early bucket (initial
        # state + LSTM weights, 62 % of the bytes) under the backward tail, the rest after the pass; buckets = 3: the side-chain bucket
        # exchanged AND updated under the main chain as well (_finish_step).  The overlapped schedules won in round 1 (tail 560 us, one
        # busy chain); with the round-2 tail (three chains that fill the GPU, 120 us shorter) NCCL's CTAs and the casts take SMs and HBM
        # from the chain that ends the step.  Measured at c2, ms per step (profiles/r2_bench_c2_n8*.json, r2_exchange_schedule.txt):
        #   8 x B200: one exchange 3.657, two buckets 3.700, three buckets 3.745;  2 x B200: one exchange 3.575, two buckets 3.593.
        self.overlap_allreduce = False
        self.buckets = 2
        self._comm_stream = None
        # grad_exchange_dtype = torch.bfloat16: the gradients cross NVLink in half the bytes (cast -> all-reduce -> Adam reads the
        # bf16 sums); every rank still applies identical updates.  Only used when world > 1.
        if grad_exchange_dtype not in (torch.float32, torch.bfloat16):
            raise TypeError("grad_exchange_dtype must be float32 or bfloat16")
        self.grad_exchange_dtype = grad_exchange_dtype
        # flat buffers; each parameter becomes a view into `flat_param` (state_dict keys/shapes unchanged)
        self.flat_param = torch.zeros(n, device=dev, dtype=torch.float32)
        self.flat_grad = torch.zeros(n, device=dev, dtype=torch.float32)
        self.exp_avg = torch.zeros(n, device=dev, dtype=torch.float32)
        self.exp_avg_sq = torch.zeros(n, device=dev, dtype=torch.float32)
        self.flat_grad16 = (torch.zeros(n, device=dev, dtype=torch.bfloat16)
                            if (self.world > 1 and grad_exchange_dtype == torch.bfloat16) else None)
        self.grad_views = {}
        for k in self.train_keys:
            p, o = named[k], offsets[k]
            m = p.numel()
            self.flat_param[o:o + m].copy_(p.data.reshape(-1))
            p.data = self.flat_param[o:o + m].view_as(p)
            self.grad_views[k] = self.flat_grad[o:o + m].view_as(p)
        self.n = off
        # optimizer state on the device: [lr, beta1, beta2, eps, weight_decay, grad_scale] and the update counter
        self.hyper = torch.tensor([self.base_lr, betas[0], betas[1], eps, 0.0, 1.0 / self.world], device=dev, dtype=torch.float32)
        self.step_dev = torch.zeros(1, device=dev, dtype=torch.int32)
        self.step_count = 0
        # StepLR (train.py:72: optim.lr_scheduler.StepLR(optimizer, args.step_size), default gamma 0.1; stepped once per
        # epoch at train.py:100)
        self.step_size, self.gamma, self.epoch = step_size, float(gamma), 0
        self.lr = self.base_lr
        self._ws = None
        self._ws_key = None
        self._a = None
        self._a_ptrs = None
        self._named = dict(decoder.named_parameters())
        self._ptr_sig = None
        self._ptr_list = None
        self._grads = None
        self._next_mask = None
        self.last_alphas = None
        self.use_cuda_graph = bool(use_cuda_graph)
        self._graphs = {}             # (feature ptr, caption ptr, shapes, training) -> (CUDAGraph, tensors kept alive)
        self._seen = {}               # the same key -> number of eager steps run so far (capture on the third)
        self.graph_replays = 0
        if self.world > 1:
            self._sync_replicas()

    # ---- data-parallel replica consistency ------------------------------------------------------------------------
    def _sync_replicas(self):
        """Every rank starts from rank 0's weights and the same flat layout.  Gradients are summed every step, so replicas
        that start equal stay equal; a replica built with a different seed / init would otherwise drift silently."""
        sig = self.layout_signature()
        sigs = [None] * self.world
        dist.all_gather_object(sigs, sig, group=self.pg)
        for r, s in enumerate(sigs):
            if s != sig:
                raise RuntimeError(f"FusedTrainStep: flat parameter layout differs between rank {dist.get_rank(self.pg)} and rank {r}")
        dist.broadcast(self.flat_param, src=dist.get_global_rank(self.pg, 0) if self.pg is not None else 0, group=self.pg)
        # parameters outside the flat buffer (frozen BERT embedding, modules the flags leave unused) are inputs too
        for k, p in self._named.items():
            if k not in self.offsets:
                dist.broadcast(p.data, src=dist.get_global_rank(self.pg, 0) if self.pg is not None else 0, group=self.pg)

    def _check_param_storage(self, cur):
        """A `dec.to()`, `.bfloat16()` or `load_state_dict(assign=True)` after construction moves a parameter out of the flat
        buffer: the forward pass would follow the new storage while Adam kept updating the old one.  Fail loudly instead."""
        base = self.flat_param.data_ptr()
        for (k, p), ptr in zip(self._named.items(), cur):
            o = self.offsets.get(k)
            if o is not None and ptr != base + 4 * o:
                raise RuntimeError(
                    f"FusedTrainStep: parameter '{k}' no longer lives in the flat parameter buffer (was the decoder moved, cast "
                    "or re-assigned after FusedTrainStep was built?).  Use load_state_dict() (copies in place) or rebuild the trainer.")

    def _args(self, img_features, captions, dropout_mask):
        """The sat_decoder_args of this step.  The struct is built once per (shape, dtype, flags) and kept; a step only
        refreshes the pointers that change (inputs, mask, stream) -- the host path between two steps is on the critical
        path of a training loop that reads the loss back every step."""
        dec = self.dec
        L = _lib.lib()
        dev = img_features.device
        B, Lloc, D = img_features.shape
        T = captions.shape[1] - 1
        a_dtype = _lib.dtype_code(img_features.dtype)
        # (strides and 16-byte alignment decide whether the library uses the features in place or compacts them into the
        # workspace, which changes the workspace size)
        key = (B, Lloc, D, T, a_dtype, bool(dec.use_tf), bool(dec.use_advanced_deep_output), bool(dec.use_attention),
               bool(dec.use_gate), dec.compute_dtype, tuple(img_features.stride()), img_features.data_ptr() % 16 == 0)
        if self._ws_key != key:
            a = _lib.DecoderArgs()
            a.struct_bytes = C.sizeof(_lib.DecoderArgs)
            a.dtype = _lib.dtype_code(dec.compute_dtype)
            a.B, a.L, a.D, a.H, a.V, a.T = B, Lloc, D, dec.embedding_size, dec.vocabulary_size, T
            a.tf, a.ado, a.attention, a.gate = (int(dec.use_tf), int(dec.use_advanced_deep_output), int(dec.use_attention),
                                                int(dec.use_gate))
            a.start_token = int(dec.start_token)
            a.need_grad, a.fused_loss, a.need_dfeat = 1, 1, 0
            a.a_dtype = a_dtype
            a.img_features = img_features.data_ptr()
            a.a_stride_b, a.a_stride_l, a.a_stride_d = img_features.stride()
            a.captions = captions.data_ptr()
            a.cap_stride = captions.stride(0)
            self.alphas = torch.empty(B, T, Lloc, device=dev, dtype=torch.float32)
            self.loss = torch.zeros(4, device=dev, dtype=torch.float32)      # [ce, alpha regulariser, total, status bits]
            # no (B,T,V) preds tensor: the fused loss reads the logits from the workspace (padded row pitch)
            a.preds, a.alphas, a.loss = None, self.alphas.data_ptr(), self.loss.data_ptr()
            nbytes = C.c_int64(0)
            _lib.check(L.sat_decoder_workspace_bytes(C.byref(a), C.byref(nbytes)), "sat_decoder_workspace_bytes")
            self._ws = torch.empty(nbytes.value + 256, device=dev, dtype=torch.uint8)
            a.workspace = self._ws.data_ptr() + ((-self._ws.data_ptr()) % 256)
            a.workspace_bytes = nbytes.value
            self._a = a
            self._a_ptrs = None           # a fresh struct: the parameter pointers have to be written again
            if self._ws_key is not None:
                self._graphs.clear()      # captured graphs point into the old workspace
                self._seen.clear()
            self._ws_key = key
        a = self._a
        a.alpha_c = float(self.alpha_c)
        a.img_features = img_features.data_ptr()
        a.a_stride_b, a.a_stride_l, a.a_stride_d = img_features.stride()
        a.captions = captions.data_ptr()
        a.cap_stride = captions.stride(0)
        a.dropout_mask = None if dropout_mask is None else dropout_mask.data_ptr()
        # parameters are views into flat_param (fixed addresses): re-read the pointers only when a storage moved
        ptrs = self._param_ptrs()
        if ptrs is not self._a_ptrs:
            for field, ptr in ptrs:
                setattr(a, field, ptr)
            self._a_ptrs = ptrs
        a.stream = torch.cuda.current_stream(dev).cuda_stream
        return a

    def _param_ptrs(self):
        """[(C field, device pointer)] of every parameter; the list object is reused while no pointer changed."""
        named = self._named
        cur = tuple(p.data_ptr() for p in named.values())
        if cur != self._ptr_sig:
            self._check_param_storage(cur)
            if self._ptr_sig is not None:     # a (non-flat) parameter moved: graphs captured its old address
                self._graphs.clear()
                self._seen.clear()
            self._ptr_sig = cur
            self._ptr_list = [(field, None if named.get(_lib.PARAM_KEYS[field]) is None else named[_lib.PARAM_KEYS[field]].data_ptr())
                              for field in _lib.PARAM_FIELDS]
        return self._ptr_list

    def _exchange(self, lo, hi):
        """All-reduce elements [lo, hi) of the flat gradient on the current stream, in the exchange dtype (async handle)."""
        if self.flat_grad16 is None:
            return dist.all_reduce(self.flat_grad[lo:hi], group=self.pg, async_op=True)
        if self.flat_grad.is_cuda:
            st = torch.cuda.current_stream(self.flat_grad.device).cuda_stream
            _lib.check(_lib.lib().sat_cast_f32_to_bf16(self.flat_grad.data_ptr() + 4 * lo, self.flat_grad16.data_ptr() + 2 * lo,
                                                       hi - lo, st), "sat_cast_f32_to_bf16")
        else:                                         # host-logic tests on CPU (gloo)
            self.flat_grad16[lo:hi].copy_(self.flat_grad[lo:hi])
        return dist.all_reduce(self.flat_grad16[lo:hi], group=self.pg, async_op=True)

    def sync_grads(self):
        """The one exchange step of data-parallel training: sum the flat gradient buffer over the ranks (NCCL over
        NVLink on GPUs).  The 1/world of the mean is folded into the Adam kernel (grad_scale)."""
        if self.world <= 1:
            return
        if not (self.overlap_allreduce and self.flat_grad.is_cuda and 0 < self.n_early < self.n):
            self._exchange(0, self.n).wait()
            return
        if self.buckets == 3:
            self._finish_step(False)
            return
        # two buckets: the early one is reduced on a side stream as soon as the backward pass has written it (under the rest
        # of the backward tail), the rest on the current stream after the pass; Adam then waits for both
        dev = self.flat_grad.device
        if self._comm_stream is None:
            self._comm_stream = torch.cuda.Stream(dev)
        cs = self._comm_stream
        _lib.check(_lib.lib().sat_wait_early_grads(cs.cuda_stream), "sat_wait_early_grads")
        with torch.cuda.stream(cs):
            early = self._exchange(0, self.n_early)
        late = self._exchange(self.n_early, self.n)
        late.wait()
        early.wait()                                   # the current stream waits for the early bucket

    def _finish_step(self, optimizer_step):
        """Exchange + update, bucket by bucket as sat_decoder_backward finishes them.  On the comm stream: the early bucket
        (initial-state and LSTM weights, final ~160 us before the pass ends) and the side-chain bucket (everything but the
        encoder projection) are all-reduced as soon as their events fire and their parameters are updated right away -- under
        the main chain's `sum_t dpre` pass and `dW_att` GEMM; on the caller's stream: the last bucket (attention.W / v, 8 % of the
        bytes at c2) after the pass, its update, the join and ONE tick of the step counter.  With world == 1 only the updates
        move (Adam's 62 us at c2 is otherwise serialised behind the pass)."""
        L = _lib.lib()
        dev = self.flat_grad.device
        if self._comm_stream is None:
            self._comm_stream = torch.cuda.Stream(dev)
        cs, cur = self._comm_stream, torch.cuda.current_stream(dev)
        nE, nS, n = self.n_early, self.n_side, self.n
        dist_on = self.world > 1

        def adam(lo, hi, stream, tick=0):
            g = self.reduced_grad()
            _lib.check(L.sat_adam_step_dev_part(self.flat_param.data_ptr() + 4 * lo, g.data_ptr() + g.element_size() * lo,
                                                _lib.dtype_code(g.dtype), self.exp_avg.data_ptr() + 4 * lo,
                                                self.exp_avg_sq.data_ptr() + 4 * lo, hi - lo, self.hyper.data_ptr(),
                                                self.step_dev.data_ptr(), tick, stream.cuda_stream), "sat_adam_step_dev_part")

        _lib.check(L.sat_wait_early_grads(cs.cuda_stream), "sat_wait_early_grads")
        with torch.cuda.stream(cs):
            if dist_on and nE > 0:
                self._exchange(0, nE).wait()
        _lib.check(L.sat_wait_side_grads(cs.cuda_stream), "sat_wait_side_grads")
        with torch.cuda.stream(cs):
            if dist_on and nS > nE:
                self._exchange(nE, nS).wait()
            if optimizer_step and nS > 0:
                adam(0, nS, cs)
        if dist_on and n > nS:
            self._exchange(nS, n).wait()
        if optimizer_step and n > nS:
            adam(nS, n, cur)
        cur.wait_stream(cs)
        if optimizer_step:
            adam(0, 0, cur, tick=1)

    def reduced_grad(self):
        """The buffer Adam consumes after sync_grads(): the fp32 flat gradient, or its all-reduced bf16 image."""
        return self.flat_grad if self.flat_grad16 is None else self.flat_grad16

    def layout_signature(self):
        """(key, offset, numel) of every trainable tensor in the flat buffers; must be identical on all ranks."""
        base = self.flat_param.data_ptr()
        named = dict(self.dec.named_parameters())
        return [(k, (named[k].data_ptr() - base) // 4, named[k].numel()) for k in self.train_keys]

    def _grad_struct(self):
        if self._grads is None:
            g = _lib.DecoderGrads()
            g.struct_bytes = C.sizeof(_lib.DecoderGrads)
            for field in _lib.PARAM_FIELDS:
                v = self.grad_views.get(_lib.PARAM_KEYS[field])
                setattr(g, field, None if v is None else v.data_ptr())
            self._grads = g
        return self._grads

    def _fresh_mask(self, B, T, device):
        """Scaled keep mask of nn.Dropout(p) (decoder.py:67): Bernoulli(1-p) / (1-p), two launches."""
        p = self.dec.dropout.p
        if not self.dec.training or p <= 0.0:
            return None
        return torch.empty(B, T, self.dec.embedding_size, device=device, dtype=torch.float32).bernoulli_(1.0 - p).mul_(1.0 / (1.0 - p))

    # ---- learning-rate schedule (train.py:72,100) --------------------------------------------------------------------
    def set_lr(self, lr):
        self.lr = float(lr)
        self.hyper[0:1].fill_(self.lr)        # a stream-ordered device write: graphs replayed afterwards read the new value

    def scheduler_step(self):
        """`scheduler.step()` of StepLR(optimizer, step_size, gamma=0.1), called once per epoch (train.py:100):
        lr = base_lr * gamma ** (epoch // step_size)."""
        self.epoch += 1
        if self.step_size:
            self.set_lr(self.base_lr * self.gamma ** (self.epoch // int(self.step_size)))
        return self.lr

    # ---- one step ------------------------------------------------------------------------------------------------------
    def _run(self, img_features, captions, dropout_mask, optimizer_step):
        L = _lib.lib()
        a = self._args(img_features, captions, dropout_mask)
        _lib.check(L.sat_decoder_forward(C.byref(a)), "sat_decoder_forward")
        _lib.check(L.sat_decoder_backward(C.byref(a), C.byref(self._grad_struct())), "sat_decoder_backward")
        if self.buckets == 3 and self.overlap_allreduce and self.flat_grad.is_cuda and 0 < self.n_early < self.n:
            self._finish_step(optimizer_step)
            return
        self.sync_grads()
        if optimizer_step:
            g = self.reduced_grad()
            fn = L.sat_adam_step_dev if g.dtype == torch.float32 else L.sat_adam_step_dev_bf16grad
            _lib.check(fn(self.flat_param.data_ptr(), g.data_ptr(), self.exp_avg.data_ptr(), self.exp_avg_sq.data_ptr(), self.n,
                          self.hyper.data_ptr(), self.step_dev.data_ptr(), a.stream), "sat_adam_step_dev")

    def step(self, img_features, captions, dropout_mask=None, optimizer_step=True):
        """One training step; returns the device tensor [ce, alpha_reg, total, status] (no host sync)."""
        dec = self.dec
        captions = dec._check_inputs(img_features, captions)      # token ids are checked lazily: check_status()
        B, T = img_features.shape[0], captions.shape[1] - 1
        with torch.cuda.device(img_features.device):
            if self.use_cuda_graph and dropout_mask is None and optimizer_step:
                self._graph_step(img_features, captions)
            else:
                own_mask = dropout_mask is None and dec.training
                if own_mask:
                    # the mask of this step was drawn at the end of the previous one (behind its Adam kernel), so a step
                    # starts with its first CUDA kernel instead of four small RNG launches
                    m = self._next_mask
                    if m is None or m.shape[0] != B or m.shape[1] != T or m.device != img_features.device:
                        m = self._fresh_mask(B, T, img_features.device)
                    dropout_mask, self._next_mask = m, None
                self._run(img_features, captions, dropout_mask, optimizer_step)
                if own_mask:
                    self._mask_in_use = dropout_mask          # keep it alive until the kernels that read it have been enqueued
                    self._next_mask = self._fresh_mask(B, T, img_features.device)
        if optimizer_step:
            self.step_count += 1
        self.last_alphas = self.alphas
        return self.loss

    def _graph_step(self, img_features, captions):
        """Replay (or, on the third step with the same buffers, capture) the CUDA graph of one whole step.  The graph is
        tied to the addresses of the feature and caption buffers: feed a training loop from a fixed set of device buffers
        (double buffering = two graphs)."""
        key = (img_features.data_ptr(), captions.data_ptr(), tuple(img_features.shape), tuple(img_features.stride()),
               img_features.dtype, tuple(captions.shape), captions.stride(0), bool(self.dec.training), float(self.alpha_c),
               float(self.dec.dropout.p), int(self.buckets))          # scalars / launch order that are baked into the captured kernel parameters
        g = self._graphs.get(key)
        if g is not None:
            g[0].replay()
            self.graph_replays += 1
            return
        B, T = img_features.shape[0], captions.shape[1] - 1
        seen = self._seen.get(key, 0)
        self._seen[key] = seen + 1
        if seen < 2 or self._ws_key is None:
            # eager steps first: they allocate the workspace, create the library's side streams / events and set the kernels'
            # shared-memory attributes -- none of which may happen during stream capture
            self._run(img_features, captions, self._fresh_mask(B, T, img_features.device), True)
            return
        if len(self._graphs) >= 8:
            self._graphs.pop(next(iter(self._graphs)))
        graph = torch.cuda.CUDAGraph()
        torch.cuda.synchronize(img_features.device)
        with torch.cuda.graph(graph):
            mask = self._fresh_mask(B, T, img_features.device)      # drawn inside the graph: a fresh mask every replay
            self._run(img_features, captions, mask, True)
        self._graphs[key] = (graph, mask, img_features, captions)
        graph.replay()              # capture records, it does not execute
        self.graph_replays += 1

    def close(self):
        """Drop the captured CUDA graphs.  Call before `dist.destroy_process_group()`: a live graph that captured NCCL all-reduces
        keeps the communicator busy and the destroy call hangs (seen with tools/dp_check.py on 2 GPUs)."""
        self._graphs.clear()
        self._seen.clear()
        if self.flat_param.is_cuda:
            torch.cuda.synchronize(self.flat_param.device)

    def check_status(self, loss_host=None):
        """Raise what the reference raises eagerly: IndexError for a caption id outside [0, V) (nn.Embedding /
        CrossEntropyLoss).  The kernels only set a flag (loss[3]); call this wherever the loss is read back anyway
        (pass the host copy to avoid another sync)."""
        flags = int((self.loss if loss_host is None else loss_host)[3])
        if flags & 1:
            raise IndexError(f"caption token out of range [0, {self.dec.vocabulary_size}) in the last step")

    def logits(self):
        """(B, T, V) fp32 view of the logits the last step left in the workspace (row pitch padded to 8 floats)."""
        if self._a is None:
            raise RuntimeError("logits() needs a step() first")
        ptr, ld = C.c_void_p(), C.c_int64()
        _lib.check(_lib.lib().sat_decoder_logits(C.byref(self._a), C.byref(ptr), C.byref(ld)), "sat_decoder_logits")
        a = self._a
        off = (ptr.value - self._ws.data_ptr()) // 4
        return torch.as_strided(self._ws.view(torch.float32), (a.B, a.T, a.V), (a.T * ld.value, ld.value, 1), off)

    def accuracy_counts(self, captions, ks=(1, 5), ignore_index=3):
        """train.py:144-145 for the step just run: device tensor [correct@1, correct@5, non-padding tokens] over
        targets = captions[:, 1:] (all T positions, utils.py:44-80), from the logits still in the workspace.  One pass over
        the logits, no host sync; accuracy in percent = 100 * correct / tokens.  ignore_index: `<pad>` = 3 in the plain
        vocabulary (generate_json_data.py:44-48), 0 for BERT."""
        from . import metrics
        if self._a is None:
            raise RuntimeError("accuracy_counts() needs a step() first")
        L = _lib.lib()
        ptr, ld = C.c_void_p(), C.c_int64()
        _lib.check(L.sat_decoder_logits(C.byref(self._a), C.byref(ptr), C.byref(ld)), "sat_decoder_logits")
        a = self._a
        targets = captions[:, 1:].to(torch.int64).contiguous()
        if targets.shape != (a.B, a.T):
            raise RuntimeError(f"captions {tuple(captions.shape)} do not belong to the last step (B={a.B}, T={a.T})")
        rank = torch.empty(a.B, a.T, device=targets.device, dtype=torch.int32)
        with torch.cuda.device(targets.device):
            _lib.check(L.sat_logit_rank(ptr.value, ld.value, a.B * a.T, a.V, targets.data_ptr(), rank.data_ptr(),
                                        torch.cuda.current_stream(targets.device).cuda_stream), "sat_logit_rank")
        return metrics.accuracy_counts(rank, targets, ks, ignore_index)
