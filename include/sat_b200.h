/* sat_b200 -- C ABI of the B200-native soft-attention LSTM caption decoder.
 *
 * Drop-in boundary for the hot path of yvokeller/Show-Attend-and-Tell (attention.py, decoder.py and the
 * loss lines of train.py).  The reference has no FFI layer: its boundary is the nn.Module surface
 * (SURVEY.md section 8(b)).  The Python modules in show-attend-and-tell_b200/ keep that surface and call
 * these entry points through ctypes.  Everything here takes plain device pointers, sizes/strides and a
 * cudaStream_t (as void*); no torch types; the library never allocates or frees device memory.
 *
 * All functions return 0 on success or a negative error code; sat_last_error() returns a thread-local
 * message.  There is no CPU fallback: every entry point requires an sm_100 device.
 */
#ifndef SAT_B200_H
#define SAT_B200_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SAT_DTYPE_F32 0   /* storage + arithmetic fp32 (true-fp32 FMA GEMMs): bit-exact beam tokens, config c1/c4 */
#define SAT_DTYPE_BF16 1  /* bf16 storage / tensor-core operands, fp32 accumulate, fp32 softmax/alphas/cell state */

/* device-side status bits of a sequence-driver call (sat_decoder_args.loss[3]) */
#define SAT_STATUS_BAD_TOKEN 1   /* a caption id outside [0, V): nn.Embedding / CrossEntropyLoss raise IndexError upstream */

/* Threading: the compute entry points may be called from several host threads (thread-local error string, per-thread / per-device
 * side streams and events, per-device kernel attributes).  The sat_set_* knobs below are DEBUG / TUNING switches stored in plain
 * process-wide ints: set them before the first compute call or from one thread while no call is in flight -- they are not part of
 * the re-entrancy guarantee, and the defaults are the product configuration. */
const char* sat_last_error(void);
int sat_version(void);
/* debug knob: 1 = run bf16 GEMMs on the SIMT kernel (bring-up cross-check of the tcgen05 kernel) */
int sat_set_force_simt(int on);
/* fp32 compute mode: 1 (default) = GEMMs whose operands are both K-major (every forward GEMM, beam search included) run on the
 * tensor cores as 3xTF32 -- each fp32 operand split into tf32 hi + lo, three tcgen05.mma.kind::tf32 per k-step into one fp32 TMEM
 * accumulator: within a few fp32 ulps of an exact-fp32 dot product; 0 = the true-fp32 SIMT FMA kernel always */
int sat_set_fp32_gemm_mode(int mode);
/* tuning knob: 1 (default) = bf16 GEMMs with K-major operands whose grid fills the GPU at least twice over (the encoder projection
 * K1) run on the 256 x 256 tile (two TMEM accumulators per CTA sharing one B stage: half the L2 traffic per flop of the 128 x 256
 * tile); 0 = always the 128-row tile.  sat_gemm's bn_override = 512 forces the wide tile (tests). */
int sat_set_gemm_wide_tile(int on);
/* debug knob: 1 = run the attention step on the generic cluster kernels even where the streaming kernels apply */
int sat_set_attn_generic(int on);
/* debug knob: work partition of the streaming attention kernels. 0 = auto, 1 = whole items per CTA, 2 = equal slot counts */
int sat_set_attn_partition(int mode);
/* debug knob: 1 (default) = PDL-aware kernels are launched with programmatic stream serialization (their prologue overlaps
 * the previous kernel's tail), 0 = ordinary launches */
int sat_set_pdl(int on);
/* tuning knob: 1 (default) = the per-step GEMMs of the decoder loop load their weight operand with the L2 evict_last hint (it
 * survives the attention kernel's 257 MB stream until the next step) and request its first tiles before the PDL wait */
int sat_set_gemm_static_weights(int on);
/* tuning knob: 1 = sat_decoder_backward runs the work after the BPTT loop as two independent chains on two streams
 * (sum over t + dW_att on the caller's stream, the small weight-gradient GEMMs / column sums on an internal side stream that
 * is joined back before the call returns), and sat_decoder_forward runs the initial-state / embedding work on that side
 * stream beside the encoder-projection GEMM; 3 (default) = like 1, and the side chain itself forks: embedding, [U; f_beta] and
 * output-layer gradients on a second internal stream beside the initial-state and LSTM weight gradients (3.61 -> 3.56 ms per
 * c2 step); 0 = one chain; 2 = like 1 with a register-thin, shared-memory-staged sum over t that GEMM CTAs of the side chain
 * can run beside (experimental: measured slower).  10 + mode (default: 13) = the main chain (sum over t, dW_att)
 * runs on an internal stream of the highest priority, so its CTAs are dispatched ahead of the side chains' pending ones (3.471 ->
 * 3.455 ms per c2 step under CUDA-graph replay; launch by launch it is 0.006 ms slower). */
int sat_set_tail_overlap(int on);
/* tuning knob: 1 (default) = with the streaming attention kernels, a backward decode step stores only its score gradients
 * de_t (B x L floats) and the sum over t of the pre-activation gradients (the input of the encoder-projection weight gradient) is
 * rebuilt after the loop in one pass over Wa; 0 = every step writes its (B, L, H) slab and one pass sums the T slabs (round-1
 * scheme: 2 x T x B x L x H x sizeof(T) more HBM traffic, T x B x L x H x sizeof(T) more workspace).  Read by
 * sat_decoder_workspace_bytes: set it before sizing the workspace. */
int sat_set_dpre_recompute(int on);
/* tuning knob: 1 (default) = with attention on, the LSTM cell backward of step t-1 runs in the epilogue of step t's dh GEMM;
 * 0 = the pointwise kernel is launched every step */
int sat_set_fused_cell_backward(int on);
/* debug instrumentation: arm (on = 1) / disarm the %globaltimer phase stamps that CTA (0,0,0) of the two big in-loop GEMMs
 * leaves (id 0 = gate GEMM, id 1 = dz GEMM; phases: 0 entry, 1 after the PDL wait, 2 first operand stage landed, 3 last MMA
 * issued, 4 accumulator complete, 5 split-K exchange done, 6 epilogue done); out (16 values) receives the latest stamps */
int sat_gemm_phase_stamps(int on, unsigned long long* out);
/* tuning knob (default 0, see DESIGN.md for the measurements): 1 = sat_decoder_forward / sat_decoder_backward run the recurrent loop of a teacher-forced batch
 * (B >= 128, streaming attention shapes) as two half-batches on two streams, so that one half's latency-bound skinny GEMMs
 * overlap the other half's HBM-bound attention kernel; 0 = whole batch on the caller's stream */
int sat_set_decoder_split(int on);
/* tuning knob of the half-batch pipelining: the attention kernel of a half uses at most attn_ctas CTAs (0 = one per item up to
 * the SM count) and the half's skinny GEMMs pick their tiles for gemm_sms SMs (0 = all) */
int sat_set_decoder_split_sms(int attn_ctas, int gemm_sms);
/* tuning knob: largest split-K cluster size a skinny GEMM may use (1 = never split, default 4, at most 8) */
int sat_set_gemm_splitk(int max_split);
/* tuning knob: L2 eviction hints of the streaming attention kernels' loads.  wa_fwd / wa_bwd: policy of the Wa rows in the
 * forward / backward kernel (0 = evict_first like the feature rows, 1 = evict_last, 2 = evict_normal); pin_bytes: additionally
 * load this many bytes of a + Wa per launch (a prefix of every item) with evict_last.  Default (1, 0, 0): the 51 MB of Wa stay
 * in the 126 MB L2 from one forward step to the next (measured -13% on the forward kernel at c2; the backward kernel is
 * fastest with everything evict_first). */
int sat_set_attn_l2_policy(int wa_fwd, int wa_bwd, int64_t pin_bytes);
/* tuning knob: the producers of the streaming attention kernels ask L2 (cp.async.bulk.prefetch.L2) for the slot they will load
 * `slots` slots later (per CTA; 0 = off), forward / backward kernel separately: bytes in flight beyond what the shared-memory
 * ring can hold. */
int sat_set_attn_prefetch(int fwd_slots, int bwd_slots);

/* Bench instrumentation: CUDA-event pairs around every attention-step (fwd = category 0, bwd = 1) and GEMM (2) launch,
 * recorded on the launching stream.  sat_profile_read fills ms[3] / counts[3] and resets. */
int sat_profile(int on);
int sat_profile_read(float* ms, int* counts);

/* train.py:144-145 / utils.py:44-80 `sequence_accuracy(preds, targets, k)`: rank[r] = position of targets[r]'s logit in row r's
 * descending order, ties broken towards the lower index (rows x V fp32 logits, row pitch ld; a target outside [0, V) or a NaN
 * target logit gets rank V).  The target is in the top k iff rank < k, so top-1 and top-5 come out of ONE pass over the logits
 * (the reference runs topk twice), and a tied block admits at most k entries like topk does. */
int sat_logit_rank(const float* logits, int64_t ld, int64_t rows, int V, const int64_t* targets, int32_t* rank, void* stream);

/* Generic GEMM  C[M,N] = act(A[M,K] * B[N,K]^T + bias[N])  (nn.Linear: attention.py:15-16, decoder.py:99,125,139-145).
 * a_kmajor: A[m*lda+k] (1) or A[k*lda+m] (0); same for B.  Cf (fp32) and/or Ct (dtype) outputs, either may be NULL. */
int sat_gemm(int dtype, int M, int N, int K, const void* A, int64_t lda, int a_kmajor, const void* B, int64_t ldb,
             int b_kmajor, const float* bias, int act, float* Cf, int64_t ldcf, void* Ct, int64_t ldct, int bn_override,
             void* stream);

/* K2: one fused attention decode step.  Replaces attention.py:14-21 (with W(img_features) hoisted, Q3) and the
 * gate decoder.py:99-100.  Rows r = img*NB + nb share image img's features (NB = 1 in training, beam width in
 * caption).  uf[r] = [U h + b_U (H) | f_beta h + b (D)] fp32.  Outputs alpha fp32, ctx / z (gated ctx) in dtype. */
int sat_attn_step_fwd(int dtype, int NI, int NB, int L, int D, int H, const void* a, const void* wa, const float* uf,
                      int64_t ld_uf, const float* v, const float* bv, int gate, float* alpha, int64_t ld_alpha,
                      void* ctx, int64_t ld_ctx, void* z, int64_t ld_z, float* ctx_f32, int64_t ld_c32, void* scratch,
                      int64_t scratch_bytes, void* stream);

/* Scratch of the streaming attention kernels (persistent CTAs over an equal split of all locations; an image cut by
 * the split leaves one partial record per piece in this buffer and a small combine kernel finishes it).  0 bytes = this
 * shape runs on the generic cluster kernels, which need none.  The buffer needs NO initialisation (every record is written
 * before it is read; tests/test_gpu_attention.py fills it with 0xFF).  Without scratch (NULL) the generic kernels run.
 * ctx_f32 (optional, streaming only) receives the fp32 un-gated context the streaming backward needs. */
int sat_attn_scratch_bytes(int NI, int NB, int L, int D, int H, int64_t* bytes);

/* No counterpart in the reference: added for callers that cannot allocate device memory
 * (code under CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc).  sat_attn_step_fwd with the scratch of
 * the streaming kernels taken from a file-scope `static __device__` array of SAT_STATIC_SCRATCH_BYTES inside the library (every GPU
 * that loads the library holds it until the process exits) instead of a caller-owned buffer; inputs and outputs are still the
 * caller's.  The call allocates nothing (no cudaMalloc / cudaFree, no allocating library routine).  Limits: only shapes the
 * streaming kernels serve (D in {512, 2048}, H in {512, 768}, L <= 1024, NB > 1 only with D = 512) and only while
 * sat_attn_scratch_bytes(NI, NB, L, D, H) <= SAT_STATIC_SCRATCH_BYTES (e.g. NI <= ~200 at D = 2048, ~1400 at D = 512, NB = 1);
 * anything else returns an error that says so.  ALL calls on a device share the one array, which keeps the previous call's
 * contents (nothing in it has to start at zero): calls that overlap on the GPU -- two streams, two host threads -- race; calls
 * ordered on one stream are safe. */
#define SAT_STATIC_SCRATCH_BYTES (4 * 1024 * 1024)
int sat_attn_step_fwd_static(int dtype, int NI, int NB, int L, int D, int H, const void* a, const void* wa, const float* uf,
                             int64_t ld_uf, const float* v, const float* bv, int gate, float* alpha, int64_t ld_alpha,
                             void* ctx, int64_t ld_ctx, void* z, int64_t ld_z, float* ctx_f32, int64_t ld_c32, void* stream);

/* K2 backward (autograd of the lines above inside loss.backward(), train.py:163).
 * The gradient wrt the hoisted projection Wa is produced in one of two ways: dpre_out != NULL -> this step's
 * dpre (B,L,H) is written once in the compute dtype (streaming kernel; the caller sums the steps afterwards);
 * dpre_out == NULL -> it is accumulated into the fp32 buffer dwa_acc (first != 0 overwrites).  The streaming kernel
 * also needs ctx_f32 (saved by sat_attn_step_fwd) and the same scratch buffer. */
int sat_attn_step_bwd(int dtype, int B, int L, int D, int H, const void* a, const void* wa, const float* uf, int64_t ld_uf,
                      const float* v, const float* alpha, int64_t ld_alpha, const void* ctx, int64_t ld_ctx,
                      const float* dz, int64_t ld_dz, const float* dctx_extra, int64_t ld_dce, const float* galpha,
                      int64_t ld_ga, int gate, void* duf, int64_t ld_duf, float* dwa_acc, float* dv_acc, float* dbv_acc,
                      float* da_acc, int first, void* dpre_out, const float* ctx_f32, int64_t ld_c32, void* scratch,
                      int64_t scratch_bytes, void* stream);

/* Arguments of the sequence drivers: Decoder.forward (decoder.py:69-135) and its backward.  Parameter pointers are
 * the fp32 tensors of the reference state_dict; NULL where a module does not exist (f_* without ado). */
typedef struct sat_decoder_args {
    int32_t struct_bytes;       /* sizeof(sat_decoder_args), checked */
    int32_t dtype;              /* SAT_DTYPE_* compute mode */
    int32_t B, L, D, H, V, T;   /* T = captions.size(1) - 1 decode steps (decoder.py:77) */
    int32_t tf, ado, attention, gate;   /* Decoder(...) flags; gate=0 is the "no gating scalar" extension (Q5) */
    int32_t start_token;        /* 0, or BERT [CLS] = 101 (decoder.py:79-82) */
    int32_t need_grad;          /* keep activations for sat_decoder_backward */
    int32_t fused_loss;         /* 1: also compute the train.py:150-162 loss (+ its gradient seed) on device */
    int32_t a_dtype;            /* dtype of img_features (SAT_DTYPE_*) */
    float alpha_c;              /* train.py:154 */
    int32_t need_dfeat;         /* 1: sat_decoder_backward will be asked for the gradient wrt img_features (train.py:132 runs the
                                 * encoder outside no_grad, so loss.backward() reaches it); sizes the workspace, set it in both calls */
    const void* img_features;   /* (B,L,D), arbitrary strides in elements (Q20) */
    int64_t a_stride_b, a_stride_l, a_stride_d;
    const int64_t* captions;    /* (B, T+1) int64, row stride cap_stride */
    int64_t cap_stride;
    const float* dropout_mask;  /* (B,T,H) fp32 scaled keep mask, NULL in eval mode (decoder.py:121-125, Q9) */
    /* parameters */
    const float* embedding;
    const float* init_h_w; const float* init_h_b; const float* init_c_w; const float* init_c_b;
    const float* f_beta_w; const float* f_beta_b;
    const float* att_U_w; const float* att_U_b; const float* att_W_w; const float* att_W_b;
    const float* att_v_w; const float* att_v_b;
    const float* lstm_w_ih; const float* lstm_w_hh; const float* lstm_b_ih; const float* lstm_b_hh;
    const float* deep_w; const float* deep_b;
    const float* f_h_w; const float* f_h_b; const float* f_z_w; const float* f_z_b; const float* f_out_w; const float* f_out_b;
    /* outputs */
    float* preds;               /* (B,T,V) fp32 (decoder.py:92); may be NULL only with fused_loss */
    float* alphas;              /* (B,T,L) fp32 (decoder.py:93) */
    float* loss;                /* [4] fp32: CE mean, alpha regulariser, total, status (SAT_FLAG_* bits as a float; 0 = clean) (fused_loss) */
    /* workspace (PyTorch-owned) */
    void* workspace;
    int64_t workspace_bytes;
    void* stream;
} sat_decoder_args;

typedef struct sat_decoder_grads {
    int32_t struct_bytes;
    int32_t reserved;
    const float* dpreds;        /* (B,T,V) fp32 upstream grad, NULL with fused_loss */
    const float* dalphas;       /* (B,T,L) fp32 upstream grad or NULL */
    const float* dloss;         /* device scalar upstream grad of the fused loss (NULL = 1) */
    /* parameter gradients, fp32, same shapes as the parameters; overwritten; NULL = skip */
    float* embedding;
    float* init_h_w; float* init_h_b; float* init_c_w; float* init_c_b;
    float* f_beta_w; float* f_beta_b;
    float* att_U_w; float* att_U_b; float* att_W_w; float* att_W_b; float* att_v_w; float* att_v_b;
    float* lstm_w_ih; float* lstm_w_hh; float* lstm_b_ih; float* lstm_b_hh;
    float* deep_w; float* deep_b;
    float* f_h_w; float* f_h_b; float* f_z_w; float* f_z_b; float* f_out_w; float* f_out_b;
    float* img_features;        /* (B,L,D) fp32 contiguous gradient wrt img_features, or NULL (needs args->need_dfeat) */
} sat_decoder_grads;

int sat_decoder_workspace_bytes(const sat_decoder_args* args, int64_t* bytes);
/* The (B, T, ld) fp32 logits the last sat_decoder_forward call with these args produced: args->preds, or the padded copy in
 * the workspace when the loss was fused (preds == NULL). */
int sat_decoder_logits(const sat_decoder_args* args, const float** logits, int64_t* ld);
/* Data-parallel overlap (no upstream equivalent; the reference is single-device): `stream` waits until the last
 * sat_decoder_backward of this host thread on the current device has written the gradients of init_h / init_c (weight, bias) and
 * lstm.weight_ih / weight_hh, which are final ~160 us before the call's last kernel; a caller can all-reduce that bucket under the
 * rest of the backward pass.  No-op before the first backward. */
int sat_wait_early_grads(void* stream);
/* ... and until every gradient except attention.W.* / attention.v.* is written (they come last, from the main chain of the
 * backward tail): a caller can all-reduce and apply all the others while the encoder projection's gradient GEMM is still running. */
int sat_wait_side_grads(void* stream);
int sat_decoder_forward(const sat_decoder_args* args);
int sat_decoder_backward(const sat_decoder_args* args, const sat_decoder_grads* grads);

/* Decoder.caption (decoder.py:160-269) batched over NI images, each with its own shrinking beam.
 * Outputs per image: n_tokens, tokens[max_len], alphas[max_len][L], status (0 = completed sentence, 1 = the
 * "No completed sentences found" fallback (Q18), in which case last_alpha holds the live rows' last-step alphas). */
typedef struct sat_beam_args {
    int32_t struct_bytes;
    int32_t dtype;
    int32_t NI, L, D, H, V;
    int32_t beam, max_steps;    /* max_steps = 50: the loop runs at most max_steps+1 iterations (Q17) */
    int32_t ado, attention, gate, bert;  /* bert selects the stop-token set {1,0} instead of {1,102} (Q15) */
    int32_t start_token;
    int32_t a_dtype;
    const void* img_features;   /* (NI,L,D) one row per image (the reference expands it to beam rows, generate_caption.py:89) */
    int64_t a_stride_b, a_stride_l, a_stride_d;
    const float* embedding;
    const float* init_h_w; const float* init_h_b; const float* init_c_w; const float* init_c_b;
    const float* f_beta_w; const float* f_beta_b;
    const float* att_U_w; const float* att_U_b; const float* att_W_w; const float* att_W_b;
    const float* att_v_w; const float* att_v_b;
    const float* lstm_w_ih; const float* lstm_w_hh; const float* lstm_b_ih; const float* lstm_b_hh;
    const float* deep_w; const float* deep_b;
    const float* f_h_w; const float* f_h_b; const float* f_z_w; const float* f_z_b; const float* f_out_w; const float* f_out_b;
    /* outputs (device) */
    int32_t* out_len;           /* (NI) number of tokens incl. start */
    int32_t* out_status;        /* (NI) */
    int64_t* out_tokens;        /* (NI, max_steps+2) */
    float* out_alphas;          /* (NI, max_steps+2, L): row 0 is all ones (decoder.py:173) */
    float* out_last_alpha;      /* (NI, beam, L) last-step alpha of the live rows (fallback return value) */
    int32_t* out_live;          /* (NI) live rows at loop exit */
    void* workspace;
    int64_t workspace_bytes;
    void* stream;
} sat_beam_args;

int sat_beam_workspace_bytes(const sat_beam_args* args, int64_t* bytes);
int sat_beam_search(const sat_beam_args* args);

/* N1 (next row): fused multi-tensor Adam over flat fp32 buffers (train.py:71,164). */
int sat_adam_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, int64_t n, float lr, float beta1,
                  float beta2, float eps, float weight_decay, int step, float grad_scale, void* stream);

/* The same update with every step-dependent scalar in DEVICE memory (CUDA-graph friendly): hyper = [lr, beta1, beta2, eps,
 * weight_decay, grad_scale] fp32; *step_count = updates done so far, incremented by the call.  StepLR (train.py:72,100) is a
 * host-side write of hyper[0] between steps. */
int sat_adam_step_dev(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, int64_t n, const float* hyper,
                      int32_t* step_count, void* stream);
/* Data-parallel gradient exchange in half the bytes (no upstream equivalent; the reference is single-device): sat_cast_f32_to_bf16
 * rounds the flat fp32 gradient to bf16, the caller all-reduces that image over NVLink, and sat_adam_step_dev_bf16grad consumes
 * the bf16 sums (same update as sat_adam_step_dev otherwise).  Every rank sees the same sums, so replicas stay identical. */
int sat_cast_f32_to_bf16(const float* src, void* dst, int64_t n, void* stream);
/* One slice [0, n) of the same update (gradient fp32 or bf16: grad_dtype = SAT_DTYPE_*), the step counter incremented only when
 * tick != 0 (n = 0: tick only): parameters whose gradients are final early are updated on another stream under the rest of the
 * backward pass, one tick at the end. */
int sat_adam_step_dev_part(float* param, const void* grad, int grad_dtype, float* exp_avg, float* exp_avg_sq, int64_t n,
                           const float* hyper, int32_t* step_count, int tick, void* stream);
int sat_adam_step_dev_bf16grad(float* param, const void* grad_bf16, float* exp_avg, float* exp_avg_sq, int64_t n, const float* hyper,
                               int32_t* step_count, void* stream);

/* N1: utils.py:101-107 `calculate_caption_lengths(captions, skip_tokens)` (train.py:174-177) without the .item(): out[0] (device
 * int64) = number of ids in the (rows, cols) grid, row pitch `stride`, equal to none of skip0..2 (-1 = unused). */
int sat_count_tokens(const int64_t* tokens, int64_t rows, int64_t cols, int64_t stride, int64_t skip0, int64_t skip1, int64_t skip2,
                     int64_t* out, void* stream);

/* N4: greedy decode of the evaluation path, train.py:272,297 `torch.max(preds, dim=2)[1]`: first maximum per row. */
int sat_argmax_rows(const float* logits, int64_t ld, int64_t rows, int V, int64_t* out, void* stream);

/* N2: HBM-resident feature cache (VGG19 is frozen, encoder.py:29-31, so train.py:132 recomputes constants every epoch):
 * out[i] = store[idx[i]] for items of item_bytes bytes (multiple of 16).  err_flag: optional device int set to 1 on an index
 * outside [0, n_items) (that item is zero-filled). */
int sat_gather_items(const void* store, int64_t item_bytes, int64_t n_items, const int64_t* idx, int64_t n_idx, void* out,
                     int32_t* err_flag, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SAT_B200_H */
