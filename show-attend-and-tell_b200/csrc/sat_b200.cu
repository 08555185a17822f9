// C-ABI entry points of libsat_b200.so (declared in include/sat_b200.h).
#include <stdarg.h>
#include "sat_decoder.cuh"
#include "sat_beam.cuh"

int g_force_simt = 0;
int g_gemm_wide = 1;         // large K-major bf16 GEMMs on the 256 x 256 two-accumulator tile (sat_gemm_tc.cuh)
int g_fp32_gemm_mode = 1;   // fp32 mode: 1 = 3xTF32 tcgen05 kernel for K-major x K-major GEMMs, 0 = SIMT FMA kernel always
int g_gemm_static_b = 1;     // in-loop GEMMs: weights loaded with evict_last and requested before the PDL wait
int g_attn_generic = 0;
int g_attn_part = 0;
int g_attn_pf_fwd = 0, g_attn_pf_bwd = 8;   // L2 prefetch distance of the streaming attention producers (slots per CTA).  Measured at c2
                                            // (tools/ab_step.py, profiles/r2_ab_attention_knobs.txt): backward 8 slots ahead = -4.5 % per
                                            // train step (bwd kernel 73 -> 66 us); any forward prefetch is slower (its Wa rows are L2-pinned)
long long g_attn_pin_bytes = 0;
int g_attn_wa_hint_fwd = 1, g_attn_wa_hint_bwd = 0;
int g_pdl = 1;
int g_dec_split = 0;      // half-batch pipelining of the recurrent loop (sat_decoder.cuh)
int g_split_attn_ctas = 0, g_split_gemm_sms = 0, g_split_chain = 1;
int g_fuse_cell_bwd = 1;  // LSTM cell backward in the epilogue of the dh GEMM (EpiDhCell)
int g_head_overlap = 1;   // forward: initial state + embeddings on the side stream beside K1
int g_beam_merge_h = 1;    // beam search: W_hh h computed by the [U; f_beta] GEMM (4H extra columns), gate GEMM without its h columns
int g_beam_presplit = 1;   // beam search, fp32 mode: per-step GEMM weights pre-split into tf32 hi / lo once per call
int g_beam_emb_table = 1;  // beam search: embedding share of the gate GEMM from a once-per-call table (sat_beam.cuh)
int g_dpre_recompute = 1;   // streaming attention backward stores de_t only; sum_t dpre_t is rebuilt in one pass over Wa (misc::dpre_sum)
int g_pack_split = 1;       // forward head: weight packing other than w_att on the side stream beside K1 (sat_set_tail_overlap(100 + ..) = off)
int g_tail_hp = 1;          // main chain of the backward tail on a highest-priority internal stream (sat_set_tail_overlap(10 + mode))
int g_tail_overlap = 3;   // backward tail as a main chain + a side chain that itself forks onto two streams (sat_decoder.cuh)
int g_gemm_max_split = 4;      // clusters of 8 launch slower than they save (tools/gemm_bench.cu)
static thread_local char g_err[1024] = "";

void sat_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

// ---- event profiling -------------------------------------------------------------------------------
#include <vector>
static int g_prof_on = 0;
static std::vector<cudaEvent_t> g_prof_ev;      // pairs (start, stop)
static std::vector<int> g_prof_cat;
static size_t g_prof_used = 0;

SatProfScope::SatProfScope(int c, cudaStream_t s) : cat(c), st(s), slot(-1) {
    if (!g_prof_on) return;
    if (g_prof_used + 2 > g_prof_ev.size()) {
        cudaEvent_t a, b;
        if (cudaEventCreate(&a) != cudaSuccess || cudaEventCreate(&b) != cudaSuccess) return;
        g_prof_ev.push_back(a); g_prof_ev.push_back(b);
    }
    slot = (int)g_prof_used;
    g_prof_used += 2;
    g_prof_cat.resize(g_prof_used / 2);
    g_prof_cat[slot / 2] = cat;
    cudaEventRecord(g_prof_ev[slot], st);
}
SatProfScope::~SatProfScope() {
    if (slot >= 0) cudaEventRecord(g_prof_ev[slot + 1], st);
}

static int require_sm100() {
    static signed char ok[64];          // per device: 0 = unknown, 1 = sm_100, -1 = something else
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) {
        sat_set_error("no CUDA device: sat_b200 has no CPU fallback");
        return SAT_ERR_CUDA;
    }
    signed char& st = ok[dev & 63];
    if (st == 0) {
        int major = 0, minor = 0;
        if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess ||
            cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev) != cudaSuccess) {
            sat_set_error("no CUDA device: sat_b200 has no CPU fallback");
            return SAT_ERR_CUDA;
        }
        st = (major == 10) ? 1 : -1;
        if (st < 0) sat_set_error("sat_b200 kernels are built for sm_100a only; found sm_%d%d", major, minor);
    }
    if (st < 0) { sat_set_error("sat_b200 kernels are built for sm_100a only"); return SAT_ERR_UNSUPPORTED; }
    return SAT_OK;
}

extern "C" {

const char* sat_last_error(void) { return g_err; }
int sat_version(void) { return 200; }
int sat_set_force_simt(int on) { g_force_simt = on ? 1 : 0; return SAT_OK; }
int sat_set_gemm_wide_tile(int on) { g_gemm_wide = on ? 1 : 0; return SAT_OK; }
int sat_set_fp32_gemm_mode(int mode) { g_fp32_gemm_mode = mode == 1 ? 1 : 0; return SAT_OK; }
int sat_set_attn_generic(int on) { g_attn_generic = on ? 1 : 0; return SAT_OK; }
int sat_set_attn_l2_policy(int wa_fwd, int wa_bwd, int64_t pin_bytes) {
    g_attn_wa_hint_fwd = (wa_fwd >= 0 && wa_fwd <= 2) ? wa_fwd : 0;
    g_attn_wa_hint_bwd = (wa_bwd >= 0 && wa_bwd <= 2) ? wa_bwd : 0;
    g_attn_pin_bytes = pin_bytes > 0 ? pin_bytes : 0;
    return SAT_OK;
}
int sat_set_gemm_splitk(int max_split) { g_gemm_max_split = max_split >= 1 ? max_split : 1; return SAT_OK; }
int sat_set_pdl(int on) { g_pdl = on ? 1 : 0; return SAT_OK; }
int sat_set_gemm_static_weights(int on) { g_gemm_static_b = on ? 1 : 0; return SAT_OK; }
int sat_set_dpre_recompute(int on) { g_dpre_recompute = on ? 1 : 0; return SAT_OK; }
int sat_set_tail_overlap(int on) { g_pack_split = on >= 100 ? 0 : 1; on %= 100; g_tail_hp = (on >= 10); on %= 10; g_tail_overlap = (on == 2 || on == 3) ? on : (on ? 1 : 0); g_head_overlap = on ? 1 : 0; return SAT_OK; }
int sat_set_fused_cell_backward(int on) { g_fuse_cell_bwd = on ? 1 : 0; return SAT_OK; }
/* debug: on = 1 arms the phase stamps of the in-loop GEMMs (gate GEMM = id 1, dz GEMM = id 2); the call also copies the last
 * stamps (nanoseconds of %globaltimer, 8 per id) into out[2][8] when out is not NULL (synchronises the device) */
int sat_gemm_phase_stamps(int on, unsigned long long* out) {
    if (out) {
        unsigned long long all[4][8];
        SAT_CHECK_CUDA(cudaDeviceSynchronize());
        SAT_CHECK_CUDA(cudaMemcpyFromSymbol(all, g_gemm_dbg, sizeof(all)));
        for (int i = 0; i < 2; ++i) for (int j = 0; j < 8; ++j) out[i * 8 + j] = all[i + 1][j];
    }
    SAT_CHECK_CUDA(cudaMemcpyToSymbol(g_gemm_dbg_on, &on, sizeof(int)));
    return SAT_OK;
}
int sat_set_decoder_split(int on) { g_dec_split = on ? 1 : 0; g_split_chain = on == 2 ? 0 : 1; return SAT_OK; }
int sat_set_decoder_split_sms(int attn_ctas, int gemm_sms) {
    g_split_attn_ctas = attn_ctas > 0 ? attn_ctas : 0; g_split_gemm_sms = gemm_sms > 0 ? gemm_sms : 0; return SAT_OK;
}
int sat_set_attn_prefetch(int fwd_slots, int bwd_slots) {
    g_attn_pf_fwd = fwd_slots > 0 ? fwd_slots : 0; g_attn_pf_bwd = bwd_slots > 0 ? bwd_slots : 0; return SAT_OK;
}
int sat_set_attn_partition(int mode) { g_attn_part = (mode == 1 || mode == 2) ? mode : 0; return SAT_OK; }

/* profiling: sat_profile(1) starts recording CUDA-event pairs around attention-step and GEMM launches on their own
 * stream; sat_profile_read() synchronises and returns per-category total ms and launch counts, then resets. */
int sat_profile(int on) { g_prof_on = on ? 1 : 0; g_prof_used = 0; return SAT_OK; }
int sat_profile_read(float* ms, int* counts) {
    for (int c = 0; c < SAT_PROF_NCAT; ++c) { ms[c] = 0.0f; counts[c] = 0; }
    for (size_t i = 0; i + 1 < g_prof_used; i += 2) {
        SAT_CHECK_CUDA(cudaEventSynchronize(g_prof_ev[i + 1]));
        float t = 0.0f;
        SAT_CHECK_CUDA(cudaEventElapsedTime(&t, g_prof_ev[i], g_prof_ev[i + 1]));
        const int c = g_prof_cat[i / 2];
        ms[c] += t; counts[c] += 1;
    }
    g_prof_used = 0;
    return SAT_OK;
}

/* Data-parallel overlap: make `stream` wait until the last sat_decoder_backward of this host thread and device has written the
 * gradients of init_h / init_c (weights and biases) and lstm.weight_ih / weight_hh -- they are final well before the call's
 * remaining work (embedding, attention, output-layer gradients) is done.  No-op before the first backward. */
int sat_wait_early_grads(void* stream) {
    SAT_TRY(require_sm100());
    SideStream* ss = nullptr;
    SAT_TRY(side_stream(&ss));
    SAT_CHECK_CUDA(cudaStreamWaitEvent(reinterpret_cast<cudaStream_t>(stream), ss->early, 0));
    return SAT_OK;
}

/* ... and until it has written every gradient except attention.W.* / attention.v.* (the end of the backward tail's side chain). */
int sat_wait_side_grads(void* stream) {
    SAT_TRY(require_sm100());
    SideStream* ss = nullptr;
    SAT_TRY(side_stream(&ss));
    SAT_CHECK_CUDA(cudaStreamWaitEvent(reinterpret_cast<cudaStream_t>(stream), ss->side, 0));
    return SAT_OK;
}

/* utils.py:44-80 (sequence_accuracy) in one pass: rank[r] = #{v : logits[r, v] > logits[r, targets[r]]} for `rows` rows of V
 * logits with row pitch ld.  The target is among the top k exactly when rank < k (a tie counts for the target). */
int sat_logit_rank(const float* logits, int64_t ld, int64_t rows, int V, const int64_t* targets, int32_t* rank, void* stream) {
    SAT_TRY(require_sm100());
    SAT_REQUIRE(logits && targets && rank && V > 0 && ld >= V && rows >= 0, "sat_logit_rank: bad arguments");
    if (rows == 0) return SAT_OK;
    SAT_REQUIRE(rows < (int64_t)1 << 31, "sat_logit_rank: too many rows");
    misc::logit_rank_kernel<<<(unsigned)rows, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(logits, ld, V, targets, rank);
    SAT_CHECK_CUDA(cudaGetLastError());
    return SAT_OK;
}

/* Where the (B, T, ld) fp32 logits of the last sat_decoder_forward with these args live: args->preds when it was given, else
 * the padded copy inside the workspace (fused loss). */
int sat_decoder_logits(const sat_decoder_args* a, const float** logits, int64_t* ld) {
    SAT_TRY(check_args(a));
    SAT_REQUIRE(logits && ld, "sat_decoder_logits: null output");
    if (a->dtype == SAT_DTYPE_F32) {
        DecWs<float> w; carve<float>(*a, w, reinterpret_cast<uint8_t*>(a->workspace));
        *logits = a->preds ? a->preds : w.logits_tmp; *ld = w.ldl;
    } else {
        DecWs<bf16> w; carve<bf16>(*a, w, reinterpret_cast<uint8_t*>(a->workspace));
        *logits = a->preds ? a->preds : w.logits_tmp; *ld = w.ldl;
    }
    return SAT_OK;
}

int sat_gemm(int dtype, int M, int N, int K, const void* A, int64_t lda, int a_kmajor, const void* B, int64_t ldb,
             int b_kmajor, const float* bias, int act, float* Cf, int64_t ldcf, void* Ct, int64_t ldct, int bn_override,
             void* stream) {
    SAT_TRY(require_sm100());
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    if (dtype == SAT_DTYPE_F32) {
        auto e = lin<float>(M, N, bias, act, Cf, ldcf, reinterpret_cast<float*>(Ct), ldct);
        return gemm<float>(st, reinterpret_cast<const float*>(A), lda, a_kmajor != 0, reinterpret_cast<const float*>(B), ldb,
                           b_kmajor != 0, M, N, K, e);
    } else if (dtype == SAT_DTYPE_BF16) {
        auto e = lin<bf16>(M, N, bias, act, Cf, ldcf, reinterpret_cast<bf16*>(Ct), ldct);
        return gemm<bf16>(st, reinterpret_cast<const bf16*>(A), lda, a_kmajor != 0, reinterpret_cast<const bf16*>(B), ldb,
                          b_kmajor != 0, M, N, K, e, bn_override);
    }
    sat_set_error("bad dtype %d", dtype);
    return SAT_ERR_ARG;
}

int sat_attn_step_fwd(int dtype, int NI, int NB, int L, int D, int H, const void* a, const void* wa, const float* uf,
                      int64_t ld_uf, const float* v, const float* bv, int gate, float* alpha, int64_t ld_alpha,
                      void* ctx, int64_t ld_ctx, void* z, int64_t ld_z, float* ctx_f32, int64_t ld_c32, void* scratch,
                      int64_t scratch_bytes, void* stream) {
    SAT_TRY(require_sm100());
    attn::FwdParams p;
    memset(&p, 0, sizeof(p));
    p.a = a; p.wa = wa; p.uf = uf; p.ld_uf = ld_uf; p.v = v; p.bv = bv; p.alpha = alpha; p.ld_alpha = ld_alpha;
    p.ctx = ctx; p.ld_ctx = ld_ctx; p.z = z; p.ld_z = ld_z;
    p.NI = NI; p.NB = NB; p.L = L; p.D = D; p.H = H; p.gate = gate;
    p.ctx32 = ctx_f32; p.ld_c32 = ld_c32;
    if (scratch && attn_s::fwd_shape_ok(NB, L, D, H) && scratch_bytes >= attn_s::scratch_bytes(NI, NB, D, H)) {
        p.s_part = reinterpret_cast<float*>(scratch);
    }
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    if (dtype == SAT_DTYPE_F32) return attn_step_fwd<float, true>(p, st);
    if (dtype == SAT_DTYPE_BF16) return attn_step_fwd<bf16, false>(p, st);
    sat_set_error("bad dtype %d", dtype);
    return SAT_ERR_ARG;
}

// ---- no-allocation variant of the attention step (requested by the run's operators; no counterpart in the reference) --------
// The streaming attention kernels need a scratch buffer for the partial records of images the work partition cuts.  Callers that
// cannot allocate (CUDA-graph capture, tight serving loops) get it from this file-scope array instead: SAT_STATIC_SCRATCH_BYTES
// of device memory on every GPU that loads the library, held until the process exits.  In host code the array's NAME is only a
// host-side stand-in, not a device address -- its device address is taken once per device with cudaGetSymbolAddress.
static __device__ __align__(256) unsigned char g_static_attn_scratch[SAT_STATIC_SCRATCH_BYTES];

static int static_attn_scratch(void** out) {
    static void* ptr[64];
    static SatPerDevice resolved;
    int dev = 0;
    SAT_CHECK_CUDA(cudaGetDevice(&dev));
    SAT_REQUIRE(dev >= 0 && dev < 64, "device index %d out of range", dev);
    SAT_TRY(resolved.once([&]() -> int {          // once per device, with that device current
        void* p = nullptr;
        SAT_CHECK_CUDA(cudaGetSymbolAddress(&p, g_static_attn_scratch));
        SAT_REQUIRE(p != nullptr && ((uintptr_t)p & 255) == 0, "static scratch: bad device address %p", p);
        ptr[dev] = p;
        return SAT_OK;
    }));
    *out = ptr[dev];
    return SAT_OK;
}

int sat_attn_step_fwd_static(int dtype, int NI, int NB, int L, int D, int H, const void* a, const void* wa, const float* uf,
                             int64_t ld_uf, const float* v, const float* bv, int gate, float* alpha, int64_t ld_alpha,
                             void* ctx, int64_t ld_ctx, void* z, int64_t ld_z, float* ctx_f32, int64_t ld_c32, void* stream) {
    SAT_TRY(require_sm100());
    SAT_REQUIRE(NI > 0 && NB > 0 && L > 0 && D > 0 && H > 0, "sat_attn_step_fwd_static: bad dims");
    SAT_REQUIRE(attn_s::fwd_shape_ok(NB, L, D, H),
                "sat_attn_step_fwd_static: shape (NB=%d, L=%d, D=%d, H=%d) is not served by the streaming kernels (D in {512, 2048}, "
                "H in {512, 768}, L <= 1024, NB > 1 only with D = 512); it needs no scratch -- call sat_attn_step_fwd with scratch = NULL",
                NB, L, D, H);
    const int64_t need = attn_s::scratch_bytes(NI, NB, D, H);
    SAT_REQUIRE(need <= (int64_t)SAT_STATIC_SCRATCH_BYTES,
                "sat_attn_step_fwd_static: NI=%d NB=%d D=%d H=%d needs %lld bytes of scratch, the static buffer holds %lld "
                "(SAT_STATIC_SCRATCH_BYTES); use sat_attn_step_fwd with a caller-owned buffer", NI, NB, D, H, (long long)need,
                (long long)SAT_STATIC_SCRATCH_BYTES);
    void* scratch = nullptr;
    SAT_TRY(static_attn_scratch(&scratch));
    // nothing in the scratch has to start at zero: every partial record is written by this launch before its combine kernel reads it
    return sat_attn_step_fwd(dtype, NI, NB, L, D, H, a, wa, uf, ld_uf, v, bv, gate, alpha, ld_alpha, ctx, ld_ctx, z, ld_z, ctx_f32,
                             ld_c32, scratch, (int64_t)SAT_STATIC_SCRATCH_BYTES, stream);
}

int sat_attn_step_bwd(int dtype, int B, int L, int D, int H, const void* a, const void* wa, const float* uf, int64_t ld_uf,
                      const float* v, const float* alpha, int64_t ld_alpha, const void* ctx, int64_t ld_ctx,
                      const float* dz, int64_t ld_dz, const float* dctx_extra, int64_t ld_dce, const float* galpha,
                      int64_t ld_ga, int gate, void* duf, int64_t ld_duf, float* dwa_acc, float* dv_acc, float* dbv_acc,
                      float* da_acc, int first, void* dpre_out, const float* ctx_f32, int64_t ld_c32, void* scratch,
                      int64_t scratch_bytes, void* stream) {
    SAT_TRY(require_sm100());
    attn::BwdParams p;
    memset(&p, 0, sizeof(p));
    p.a = a; p.wa = wa; p.uf = uf; p.ld_uf = ld_uf; p.v = v; p.alpha = alpha; p.ld_alpha = ld_alpha;
    p.ctx = ctx; p.ld_ctx = ld_ctx; p.dz = dz; p.ld_dz = ld_dz; p.dctx_extra = dctx_extra; p.ld_dce = ld_dce;
    p.galpha = galpha; p.ld_ga = ld_ga; p.duf = duf; p.ld_duf = ld_duf; p.dwa_acc = dwa_acc; p.dv_acc = dv_acc;
    p.dbv_acc = dbv_acc; p.da_acc = da_acc; p.first = first; p.dpre_out = dpre_out;
    p.B = B; p.L = L; p.D = D; p.H = H; p.gate = gate;
    p.ctx32 = ctx_f32; p.ld_c32 = ld_c32;
    if (scratch && attn_s::shape_ok(L, D, H) && scratch_bytes >= attn_s::scratch_bytes(B, 1, D, H)) {
        p.s_part = reinterpret_cast<float*>(scratch);
        p.s_dctx = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(scratch) + attn_s::scratch_rec_bytes(1, D, H));
    }
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    if (dtype == SAT_DTYPE_F32) return attn_step_bwd<float, true>(p, st);
    if (dtype == SAT_DTYPE_BF16) return attn_step_bwd<bf16, false>(p, st);
    sat_set_error("bad dtype %d", dtype);
    return SAT_ERR_ARG;
}

int sat_attn_scratch_bytes(int NI, int NB, int L, int D, int H, int64_t* bytes) {
    SAT_REQUIRE(bytes != nullptr, "null out pointer");
    SAT_TRY(require_sm100());
    *bytes = attn_s::fwd_shape_ok(NB, L, D, H) ? attn_s::scratch_bytes(NI, NB, D, H) : 0;
    return SAT_OK;
}

int sat_decoder_workspace_bytes(const sat_decoder_args* args, int64_t* bytes) {
    SAT_TRY(check_args(args));
    SAT_REQUIRE(bytes != nullptr, "null out pointer");
    if (args->dtype == SAT_DTYPE_F32) { DecWs<float> w; *bytes = carve<float>(*args, w, nullptr); }
    else { DecWs<bf16> w; *bytes = carve<bf16>(*args, w, nullptr); }
    return SAT_OK;
}

int sat_decoder_forward(const sat_decoder_args* args) {
    SAT_TRY(check_args(args));
    SAT_TRY(require_sm100());
    return args->dtype == SAT_DTYPE_F32 ? decoder_forward_t<float>(*args) : decoder_forward_t<bf16>(*args);
}

int sat_decoder_backward(const sat_decoder_args* args, const sat_decoder_grads* grads) {
    SAT_TRY(check_args(args));
    SAT_TRY(require_sm100());
    SAT_REQUIRE(grads && grads->struct_bytes == (int32_t)sizeof(sat_decoder_grads), "sat_decoder_grads size mismatch");
    return args->dtype == SAT_DTYPE_F32 ? decoder_backward_t<float>(*args, *grads) : decoder_backward_t<bf16>(*args, *grads);
}

int sat_beam_workspace_bytes(const sat_beam_args* args, int64_t* bytes) {
    SAT_TRY(beam_check_args(args));
    SAT_REQUIRE(bytes != nullptr, "null out pointer");
    if (args->dtype == SAT_DTYPE_F32) { BeamWs<float> w; *bytes = beam_carve<float>(*args, w, nullptr); }
    else { BeamWs<bf16> w; *bytes = beam_carve<bf16>(*args, w, nullptr); }
    return SAT_OK;
}

int sat_beam_search(const sat_beam_args* args) {
    SAT_TRY(beam_check_args(args));
    SAT_TRY(require_sm100());
    return args->dtype == SAT_DTYPE_F32 ? beam_search_t<float>(*args) : beam_search_t<bf16>(*args);
}

__global__ void adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                            int64_t n, float lr, float b1, float b2, float eps, float wd, float bc1, float bc2, float gs) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        float gr = g[i] * gs;
        const float pv = p[i];
        if (wd != 0.0f) gr += wd * pv;
        const float mi = b1 * m[i] + (1.0f - b1) * gr;
        const float vi = b2 * v[i] + (1.0f - b2) * gr * gr;
        m[i] = mi; v[i] = vi;
        const float denom = sqrtf(vi) / sqrtf(bc2) + eps;       // torch.optim.Adam (train.py:71)
        p[i] = pv - (lr / bc1) * (mi / denom);
    }
}

int sat_adam_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, int64_t n, float lr, float beta1,
                  float beta2, float eps, float weight_decay, int step, float grad_scale, void* stream) {
    SAT_TRY(require_sm100());
    if (n <= 0) return SAT_OK;
    const float bc1 = 1.0f - powf(beta1, (float)step), bc2 = 1.0f - powf(beta2, (float)step);
    adam_kernel<<<misc::grid_for(n), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(param, grad, exp_avg, exp_avg_sq, n, lr, beta1,
                                                                                      beta2, eps, weight_decay, bc1, bc2, grad_scale);
    SAT_CHECK_CUDA(cudaGetLastError());
    return SAT_OK;
}

/* Same update with the step-dependent scalars read from DEVICE memory, so that a whole train step can live in a CUDA graph:
 * hyper = [lr, beta1, beta2, eps, weight_decay, grad_scale]; *step_count is the number of updates done so far (the kernel uses
 * step_count + 1 for the bias corrections, a one-thread kernel increments it afterwards).  StepLR (train.py:72,100) then is a
 * host-side write of hyper[0] between steps. */
}  // extern "C"  (the kernel below is a template)
template <typename TG>
__global__ void adam_dev_kernel(float* __restrict__ p, const TG* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                                int64_t n, const float* __restrict__ hyper, const int32_t* __restrict__ step_count) {
    const float lr = hyper[0], b1 = hyper[1], b2 = hyper[2], eps = hyper[3], wd = hyper[4], gs = hyper[5];
    const float step = (float)(step_count[0] + 1);
    const float bc1 = 1.0f - powf(b1, step), bc2 = 1.0f - powf(b2, step);
    const float inv_sbc2 = 1.0f / sqrtf(bc2), slr = lr / bc1;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        float gr = to_f32<TG>(g[i]) * gs;
        const float pv = p[i];
        if (wd != 0.0f) gr += wd * pv;
        const float mi = b1 * m[i] + (1.0f - b1) * gr;
        const float vi = b2 * v[i] + (1.0f - b2) * gr * gr;
        m[i] = mi; v[i] = vi;
        p[i] = pv - slr * (mi / (sqrtf(vi) * inv_sbc2 + eps));
    }
}
__global__ void adam_tick_kernel(int32_t* step_count) { step_count[0] += 1; }
extern "C" {

static int adam_step_dev_any(float* param, const void* grad, int grad_dtype, float* exp_avg, float* exp_avg_sq, int64_t n,
                             const float* hyper, int32_t* step_count, void* stream) {
    SAT_TRY(require_sm100());
    SAT_REQUIRE(hyper && step_count, "sat_adam_step_dev: null hyper / step_count");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    if (n > 0) {
        if (grad_dtype == SAT_DTYPE_F32)
            adam_dev_kernel<float><<<misc::grid_for(n), 256, 0, st>>>(param, reinterpret_cast<const float*>(grad), exp_avg, exp_avg_sq, n, hyper, step_count);
        else
            adam_dev_kernel<bf16><<<misc::grid_for(n), 256, 0, st>>>(param, reinterpret_cast<const bf16*>(grad), exp_avg, exp_avg_sq, n, hyper, step_count);
    }
    adam_tick_kernel<<<1, 1, 0, st>>>(step_count);
    SAT_CHECK_CUDA(cudaGetLastError());
    return SAT_OK;
}
int sat_adam_step_dev(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, int64_t n, const float* hyper,
                      int32_t* step_count, void* stream) {
    return adam_step_dev_any(param, grad, SAT_DTYPE_F32, exp_avg, exp_avg_sq, n, hyper, step_count, stream);
}
/* ... with the gradient in bf16: what a data-parallel step all-reduced in half the bytes (sat_cast_f32_to_bf16) */
int sat_adam_step_dev_bf16grad(float* param, const void* grad_bf16, float* exp_avg, float* exp_avg_sq, int64_t n, const float* hyper,
                               int32_t* step_count, void* stream) {
    return adam_step_dev_any(param, grad_bf16, SAT_DTYPE_BF16, exp_avg, exp_avg_sq, n, hyper, step_count, stream);
}
/* One slice of the update: the same arithmetic on elements [0, n) of the given pointers, gradient fp32 or bf16 (grad_dtype =
 * SAT_DTYPE_*); the step counter is incremented only when tick != 0 (n may be 0: tick only).  Lets a caller update the parameters
 * whose gradients are final early on another stream under the rest of the backward pass, and tick once at the end. */
int sat_adam_step_dev_part(float* param, const void* grad, int grad_dtype, float* exp_avg, float* exp_avg_sq, int64_t n,
                           const float* hyper, int32_t* step_count, int tick, void* stream) {
    SAT_TRY(require_sm100());
    SAT_REQUIRE(hyper && step_count, "sat_adam_step_dev_part: null hyper / step_count");
    SAT_REQUIRE(grad_dtype == SAT_DTYPE_F32 || grad_dtype == SAT_DTYPE_BF16, "sat_adam_step_dev_part: bad grad dtype %d", grad_dtype);
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    if (n > 0) {
        if (grad_dtype == SAT_DTYPE_F32)
            adam_dev_kernel<float><<<misc::grid_for(n), 256, 0, st>>>(param, reinterpret_cast<const float*>(grad), exp_avg, exp_avg_sq, n, hyper, step_count);
        else
            adam_dev_kernel<bf16><<<misc::grid_for(n), 256, 0, st>>>(param, reinterpret_cast<const bf16*>(grad), exp_avg, exp_avg_sq, n, hyper, step_count);
    }
    if (tick) adam_tick_kernel<<<1, 1, 0, st>>>(step_count);
    SAT_CHECK_CUDA(cudaGetLastError());
    return SAT_OK;
}

/* Gradient compression for the data-parallel exchange: dst (bf16) = src (fp32), n elements. */
int sat_cast_f32_to_bf16(const float* src, void* dst, int64_t n, void* stream) {
    SAT_TRY(require_sm100());
    SAT_REQUIRE((src && dst) || n == 0, "sat_cast_f32_to_bf16: null pointer");
    return misc::cast<float, bf16>(reinterpret_cast<cudaStream_t>(stream), reinterpret_cast<bf16*>(dst), src, n);
}

/* utils.py:101-107 / train.py:174-177: tokens of the (rows, cols) int64 grid (row pitch `stride`) that are none of the three
 * skip ids (pass -1 for an unused one); the int64 count is written to out[0] on the device. */
int sat_count_tokens(const int64_t* tokens, int64_t rows, int64_t cols, int64_t stride, int64_t skip0, int64_t skip1, int64_t skip2,
                     int64_t* out, void* stream) {
    SAT_TRY(require_sm100());
    SAT_REQUIRE(tokens && out && rows >= 0 && cols >= 0, "sat_count_tokens: bad arguments");
    misc::count_tokens_kernel<<<1, 1024, 0, reinterpret_cast<cudaStream_t>(stream)>>>(tokens, rows, cols, stride, skip0, skip1, skip2, out);
    SAT_CHECK_CUDA(cudaGetLastError());
    return SAT_OK;
}

/* train.py:272,297 `torch.max(preds, dim=2)[1]`: first maximum of every row of (rows, V) fp32 logits with row pitch ld. */
int sat_argmax_rows(const float* logits, int64_t ld, int64_t rows, int V, int64_t* out, void* stream) {
    SAT_TRY(require_sm100());
    SAT_REQUIRE(logits && out && V > 0 && ld >= V && rows >= 0 && rows < ((int64_t)1 << 31), "sat_argmax_rows: bad arguments");
    if (rows == 0) return SAT_OK;
    misc::argmax_rows_kernel<<<(unsigned)rows, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(logits, ld, V, out, 1);
    SAT_CHECK_CUDA(cudaGetLastError());
    return SAT_OK;
}

/* N2 feature cache: out[i] = store[idx[i]] for items of item_bytes bytes (a multiple of 16; both buffers 16-byte aligned).
 * err_flag (device int, optional) receives 1 if an index is outside [0, n_items); that item is zero-filled. */
int sat_gather_items(const void* store, int64_t item_bytes, int64_t n_items, const int64_t* idx, int64_t n_idx, void* out,
                     int32_t* err_flag, void* stream) {
    SAT_TRY(require_sm100());
    SAT_REQUIRE(store && idx && out && item_bytes > 0 && item_bytes % 16 == 0 && n_items > 0 && n_idx >= 0, "sat_gather_items: bad arguments");
    SAT_REQUIRE(al16(store) && al16(out), "sat_gather_items: buffers must be 16-byte aligned");
    if (n_idx == 0) return SAT_OK;
    SAT_REQUIRE(n_idx < 65536, "sat_gather_items: at most 65535 indices per call");
    const int64_t rv = item_bytes / 16;
    int bx = (int)(sat_cdiv64(rv, 256 * 4));          // ~4 vectors per thread
    const int64_t cap = sat_cdiv64((int64_t)148 * 16, n_idx);
    if (bx > cap) bx = (int)cap;
    if (bx < 1) bx = 1;
    dim3 grid(bx, (unsigned)n_idx);
    misc::gather_items_kernel<<<grid, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        reinterpret_cast<const uint4*>(store), rv, n_items, idx, reinterpret_cast<uint4*>(out), err_flag);
    SAT_CHECK_CUDA(cudaGetLastError());
    return SAT_OK;
}

}  // extern "C"
