// Sequence drivers: Decoder.forward (decoder.py:69-135) and its backward (loss.backward(), train.py:163).
// Host code only launches kernels on the caller's stream; every buffer lives in the PyTorch-owned workspace.
#pragma once
#include <type_traits>
#include "../../include/sat_b200.h"
#include "sat_common.cuh"
#include "sat_epilogue.cuh"
#include "sat_gemm_simt.cuh"
#include "sat_gemm_tc.cuh"
#include "sat_gemm_tf32.cuh"
#include "sat_attn_stream.cuh"
#include "sat_misc.cuh"

extern int g_force_simt, g_gemm_static_b;
// phase-stamp ids of the two big in-loop GEMMs (debug instrumentation in sat_gemm_tc.cuh)
template <typename T, bool P> struct EpiDbgId<EpiLstm<T, P>> { static constexpr int value = 1; };
template <typename T, bool P> struct EpiDbgId<EpiDzGate<T, P>> { static constexpr int value = 2; };
// ... and both drain their accumulator through the coalesced path (sat_gemm_tc.cuh)
template <typename T, bool P> struct EpiCoalesce<EpiLstm<T, P>> { static constexpr bool value = true; };
template <typename T, bool P> struct EpiCoalesce<EpiDzGate<T, P>> { static constexpr bool value = true; };
template <typename T, bool P> struct EpiCoalesce<EpiDhCell<T, P>> { static constexpr bool value = true; };
// (EpiLinear too: measured on the whole c2 step 4.00 -> 3.94 ms, K1 + Xg + vocabulary GEMMs 238 -> 216 us)
template <typename T, bool P> struct EpiCoalesce<EpiLinear<T, P>> { static constexpr bool value = true; };

template <typename T> struct IsF32 { static constexpr bool value = std::is_same<T, float>::value; };

template <typename T, typename Epi>
int gemm(cudaStream_t s, const T* A, int64_t lda, bool a_k, const T* B, int64_t ldb, bool b_k, int M, int N, int K,
         const Epi& epi, int bn = 0, int sm_budget = 0, int b_static = 0, const T* B_hi = nullptr, const T* B_lo = nullptr) {
    SatProfScope prof(SAT_PROF_GEMM, s);
    if constexpr (IsF32<T>::value) {
        // fp32 mode: forward-shaped GEMMs (both operands K-major) run on the tensor cores as 3xTF32 (fp32-accurate, see
        // sat_gemm_tf32.cuh); transposed-operand (gradient) GEMMs and tiny ones stay on the true-fp32 SIMT kernel
        // (B_hi, B_lo): the same operand pre-split by tc32::split_operand (fp32 mode only; the SIMT fallback reads the original B)
        if (tc32::applies(A, lda, a_k, B, ldb, b_k, M, N, K)) {
            if (B_hi && B_lo && ((((uintptr_t)B_hi | (uintptr_t)B_lo) & 15) == 0)) return tc32::launch<Epi>(s, A, lda, B_hi, ldb, M, N, K, epi, B_lo);
            return tc32::launch<Epi>(s, A, lda, B, ldb, M, N, K, epi);
        }
        return simt::launch<float, Epi>(s, A, lda, a_k, B, ldb, b_k, M, N, K, epi);
    } else {
        if (g_force_simt) return simt::launch<bf16, Epi>(s, A, lda, a_k, B, ldb, b_k, M, N, K, epi);
        return tc::launch<Epi>(s, A, lda, a_k, B, ldb, b_k, M, N, K, epi, bn, 0, sm_budget, g_gemm_static_b ? b_static : 0);
    }
}

static inline bool al16(const void* p) { return ((uintptr_t)p & 15) == 0; }

// ---------------------------------------------------------------------------------------------------
// Half-batch pipelining of the recurrent loop.  The decode steps of different captions never exchange data, so the
// batch is cut into two halves that run the whole T-step loop on two streams: while one half's attention kernel
// streams a / Wa out of HBM (memory-bound, one CTA per item), the other half's latency-bound skinny GEMMs (U h, gates,
// dz, dh) run on the SMs the attention grid leaves free.  Everything outside the loop still sees the full (t, b) row
// space; a half is only a row range [b0, b0 + nb) of every time-major slice.  The side stream and its two events are
// created once per host thread and device; the caller's stream forks into it before the loop and joins after it.
// ---------------------------------------------------------------------------------------------------
extern int g_dec_split, g_split_attn_ctas, g_split_gemm_sms, g_split_chain, g_tail_overlap, g_fuse_cell_bwd, g_head_overlap, g_dpre_recompute, g_tail_hp, g_pack_split;
struct SideStream { cudaStream_t s, s2, hp; cudaEvent_t fork, join, join2, joinhp, attn_done[2], early, side; bool ok; };
static inline int side_stream(SideStream** out) {
    static thread_local SideStream cache[32] = {};
    int dev = 0;
    SAT_CHECK_CUDA(cudaGetDevice(&dev));
    SAT_REQUIRE(dev >= 0 && dev < 32, "device index %d out of range", dev);
    SideStream& ss = cache[dev];
    if (!ss.ok) {
        SAT_CHECK_CUDA(cudaStreamCreateWithFlags(&ss.s, cudaStreamNonBlocking));
        SAT_CHECK_CUDA(cudaStreamCreateWithFlags(&ss.s2, cudaStreamNonBlocking));
        {   // the critical chain of the backward tail may run on a stream of the highest priority (g_tail_hp): its CTAs are
            // dispatched ahead of the side chains' pending ones whenever an SM frees up
            int least = 0, greatest = 0;
            SAT_CHECK_CUDA(cudaDeviceGetStreamPriorityRange(&least, &greatest));
            SAT_CHECK_CUDA(cudaStreamCreateWithPriority(&ss.hp, cudaStreamNonBlocking, greatest));
        }
        SAT_CHECK_CUDA(cudaEventCreateWithFlags(&ss.joinhp, cudaEventDisableTiming));
        SAT_CHECK_CUDA(cudaEventCreateWithFlags(&ss.fork, cudaEventDisableTiming));
        SAT_CHECK_CUDA(cudaEventCreateWithFlags(&ss.join, cudaEventDisableTiming));
        SAT_CHECK_CUDA(cudaEventCreateWithFlags(&ss.join2, cudaEventDisableTiming));
        SAT_CHECK_CUDA(cudaEventCreateWithFlags(&ss.attn_done[0], cudaEventDisableTiming));
        SAT_CHECK_CUDA(cudaEventCreateWithFlags(&ss.attn_done[1], cudaEventDisableTiming));
        SAT_CHECK_CUDA(cudaEventCreateWithFlags(&ss.early, cudaEventDisableTiming));
        SAT_CHECK_CUDA(cudaEventCreateWithFlags(&ss.side, cudaEventDisableTiming));
        ss.ok = true;
    }
    *out = &ss;
    return SAT_OK;
}

template <typename T>
EpiLinear<T, IsF32<T>::value> lin(int M, int N, const float* bias, int act, float* Cf, int64_t ldcf, T* Ct, int64_t ldct,
                                  const float* addend = nullptr, int64_t ld_add = 0) {
    EpiLinear<T, IsF32<T>::value> e;
    e.bias = bias; e.addend = addend; e.ld_add = ld_add; e.Cf = Cf; e.ldcf = ldcf; e.Ct = Ct; e.ldct = ldct;
    e.mulmask = nullptr; e.ld_mm = 0; e.act = act; e.M = M; e.N = N; e.remapB = 0; e.remapT = 0;
    e.vec_ok = al16(bias) && al16(addend) && al16(Cf) && al16(Ct) && (ld_add % 8 == 0) && (ldcf % 8 == 0) && (ldct % 8 == 0);
    return e;
}

// ---------------------------------------------------------------------------------------------------
// workspace carve-up (identical in forward and backward).  Internal activations are TIME-MAJOR (t, b, .)
// so that every per-step slice is a contiguous (B x width) matrix and the batched weight-gradient GEMMs
// contract over one affine row index k = t*B + b.  Reference-facing tensors keep the reference layouts.
// ---------------------------------------------------------------------------------------------------
template <typename T>
struct DecWs {
    T *w_att, *w_uf, *w_gx, *w_gzh, *w_init, *w_out, *w_fh, *w_fz;     // packed compute-dtype weights
    float *b_uf, *b_g, *b_init;
    T *a_c, *wa, *abar;
    float* abar_f;
    float* mean_part; // (4,B,D) partial sums of the mean over L
    T* E;            // (T,B,H)     embeddings fed to each step
    float* Xg;       // (T,B,4H)    embedding share of the gate pre-activations (+ biases), interleaved
    float* UF;       // (T,B,H+D)   [U h | f_beta h] pre-activations
    T* ZH;           // (T+1,B,D+H) [gated ctx z_t | h_t]  (h_t = state entering step t)
    float* Cst;      // (T+1,B,H)   cell states
    T* Gates;        // (T,B,4H)    activated gates
    T* Ctx;          // (T,B,D)     un-gated context
    T* Hdrop;        // (T,B,H)     dropout(h_{t+1})
    float* hc0;      // (B,2H)      [h0 | c0]
    T *P1, *P2, *Comb;   // ado (T,B,H)
    int64_t* tok_in; // (T,B) token fed to each step (free-running mode)
    int* flags;      // [4] device-side status bits of the call (SAT_FLAG_*)
    float* rowloss;  // (B*T)
    float* loss_part; // (2 * LOSS_PARTS) partial sums of the loss reduction
    float* logits_tmp;   // (B,T,Vp) when preds == NULL: row pitch padded to 8 so the GEMM epilogue stores 16-byte vectors
    int64_t ldl;         // row pitch of the logits the drivers work on: V (preds) or Vp (logits_tmp)
    // backward
    T* dlog;         // (T,B,Vp)
    float* dHd;      // (T,B,H)
    T* dG;           // (T,B,4H)
    float* dZH;      // (B,D+H), or (T,B,D+H) when the gradient wrt img_features is wanted (every step's dctx is kept)
    float* dabar;    // (B,D)  gradient wrt the mean feature (need_dfeat)
    T* dUF;          // (T,B,H+D)
    float *dh_carry, *dc_carry;
    float* dwa_acc;  // (B,L,H)      generic attention backward: fp32 accumulator over time
    T* dpre_all;     // (T,B,L,H)    streaming attention backward: each step's dpre written once (g_dpre_recompute == 0)
    float* de_all;   // (T,B,L)      streaming attention backward: each step's score gradients (g_dpre_recompute == 1)
    float* Ctx32;    // (T,B,D)      fp32 un-gated context saved by the streaming forward kernel for its backward
    uint8_t* attn_scratch;   // streaming attention kernels: arrival counters + partial records
    uint8_t* attn_scratch2;
    int64_t attn_scratch_bytes;
    bool attn_stream;
    T* dwa_t;        // (B,L,H)
    float *dv_acc, *dbv_acc;
    T* dE;           // (T,B,H)
    T* dpre0;        // (B,2H)
    float* galpha;   // (B,T,L)
    float* gscratch;
    float* csum;
    float* csum2;
    float* csum3;
    float* gscratch2;
    float* dcomb;    // ado (T,B,H) fp32
    T *dP1, *dP2;    // ado (T,B,H)
    float* dctx_all; // ado (T,B,D) fp32
    int64_t csum_floats;
    int64_t Vp;
    int64_t total;
};

template <typename T>
int64_t carve(const sat_decoder_args& a, DecWs<T>& w, uint8_t* base) {
    int64_t off = 0;
    auto take = [&](int64_t bytes) -> void* {
        void* p = base ? base + off : nullptr;
        off += (bytes + 255) & ~(int64_t)255;
        return p;
    };
    const int64_t B = a.B, L = a.L, D = a.D, H = a.H, V = a.V, Tn = a.T;
    w.Vp = (V + 7) & ~(int64_t)7;
#define TK(field, count, type) w.field = reinterpret_cast<type*>(take((int64_t)(count) * (int64_t)sizeof(type)))
    TK(w_att, H * D, T); TK(w_uf, (H + D) * H, T); TK(w_gx, 4 * H * H, T); TK(w_gzh, 4 * H * (D + H), T);
    TK(w_init, 2 * H * D, T); TK(w_out, V * H, T);
    TK(w_fh, a.ado ? H * H : 0, T); TK(w_fz, a.ado ? H * D : 0, T);
    TK(b_uf, H + D, float); TK(b_g, 4 * H, float); TK(b_init, 2 * H, float);
    const bool a_alias = (a.a_dtype == a.dtype) && a.a_stride_d == 1 && a.a_stride_l == D && a.a_stride_b == L * D &&
                         al16(a.img_features);
    TK(a_c, a_alias ? 0 : B * L * D, T);
    if (a_alias) w.a_c = reinterpret_cast<T*>(const_cast<void*>(a.img_features));
    TK(wa, a.attention ? B * L * H : 0, T);
    w.attn_stream = a.attention && attn_s::fwd_shape_ok(1, a.L, a.D, a.H);
    w.attn_scratch_bytes = w.attn_stream ? attn_s::scratch_bytes(a.B, 1, a.D, a.H) : 0;
    TK(attn_scratch, w.attn_scratch_bytes, uint8_t);
    TK(attn_scratch2, w.attn_scratch_bytes, uint8_t);      // second half of a split batch (half-batch pipelining)
    TK(Ctx32, (w.attn_stream && a.need_grad) ? B * Tn * D : 0, float);
    TK(abar, B * D, T); TK(abar_f, B * D, float); TK(mean_part, 4 * B * D, float);
    TK(E, B * Tn * H, T); TK(Xg, B * Tn * 4 * H, float); TK(UF, B * Tn * (H + D), float);
    TK(ZH, B * (Tn + 1) * (D + H), T); TK(Cst, B * (Tn + 1) * H, float);
    TK(Gates, B * Tn * 4 * H, T); TK(Ctx, B * Tn * D, T); TK(Hdrop, B * Tn * H, T); TK(hc0, B * 2 * H, float);
    TK(P1, a.ado ? B * Tn * H : 0, T); TK(P2, a.ado ? B * Tn * H : 0, T); TK(Comb, a.ado ? B * Tn * H : 0, T);
    TK(tok_in, B * Tn, int64_t); TK(flags, 4, int); TK(rowloss, B * Tn, float); TK(loss_part, 2 * misc::LOSS_PARTS, float);
    TK(logits_tmp, a.preds ? 0 : B * Tn * w.Vp, float);
    w.ldl = a.preds ? (int64_t)V : w.Vp;
    if (a.need_grad) {
        TK(dlog, B * Tn * w.Vp, T); TK(dHd, B * Tn * H, float); TK(dG, B * Tn * 4 * H, T); TK(dZH, (a.need_dfeat ? Tn : 1) * B * (D + H), float);
        TK(dabar, a.need_dfeat ? B * D : 0, float);
        TK(dUF, B * Tn * (H + D), T); TK(dh_carry, B * H, float); TK(dc_carry, B * H, float);
        TK(dwa_acc, (a.attention && !w.attn_stream) ? B * L * H : 0, float); TK(dwa_t, a.attention ? B * L * H : 0, T);
        // streaming attention: the T slabs of dpre (g_dpre_recompute == 0), or only the score gradients de_t they are rebuilt from
        TK(dpre_all, (w.attn_stream && !g_dpre_recompute) ? Tn * B * L * H : 0, T);
        TK(de_all, (w.attn_stream && g_dpre_recompute) ? Tn * B * L : 0, float);
        if (!(w.attn_stream && g_dpre_recompute)) w.de_all = nullptr;      // (a zero-byte take still hands out an address)
        else w.dpre_all = nullptr;
        TK(dv_acc, B * H, float); TK(dbv_acc, B, float);
        TK(dE, B * Tn * H, T); TK(dpre0, B * 2 * H, T); TK(galpha, B * Tn * L, float);
        int64_t gs = 4 * H * (D + H);
        if (gs < (H + D) * H) gs = (H + D) * H;
        if (gs < 2 * H * D) gs = 2 * H * D;
        TK(gscratch, gs, float);
        w.csum_floats = 64 * (V > 4 * H ? V : 4 * H);
        if (w.csum_floats < 64 * (H + D)) w.csum_floats = 64 * (H + D);
        if (w.csum_floats < (int64_t)4 * 148 * H) w.csum_floats = (int64_t)4 * 148 * H;      // sum_over_t_colsum: one row per block
        TK(csum, w.csum_floats, float);
        TK(csum2, w.csum_floats, float);      // column-sum scratch of the side chain of the backward tail
        TK(csum3, w.csum_floats, float);      // ... and of its second branch
        TK(gscratch2, (H + D) * H, float);
        TK(dcomb, a.ado ? B * Tn * H : 0, float); TK(dP1, a.ado ? B * Tn * H : 0, T); TK(dP2, a.ado ? B * Tn * H : 0, T);
        TK(dctx_all, a.ado ? B * Tn * D : 0, float);
    }
#undef TK
    w.total = off;
    return off;
}

static inline int check_args(const sat_decoder_args* a) {
    SAT_REQUIRE(a != nullptr, "null args");
    SAT_REQUIRE(a->struct_bytes == (int32_t)sizeof(sat_decoder_args), "sat_decoder_args size mismatch: got %d, library expects %zu",
                a->struct_bytes, sizeof(sat_decoder_args));
    SAT_REQUIRE(a->dtype == SAT_DTYPE_F32 || a->dtype == SAT_DTYPE_BF16, "bad dtype %d", a->dtype);
    SAT_REQUIRE(a->B > 0 && a->L > 0 && a->D > 0 && a->H > 0 && a->V > 0 && a->T > 0, "bad dims");
    SAT_REQUIRE(a->D % 8 == 0 && a->H % 8 == 0, "D (%d) and H (%d) must be multiples of 8", a->D, a->H);
    SAT_REQUIRE(a->alphas != nullptr && (a->preds != nullptr || a->fused_loss), "missing output buffers");
    return SAT_OK;
}

// ---------------------------------------------------------------------------------------------------
// forward
// ---------------------------------------------------------------------------------------------------
// `rest`: stream of everything but the encoder projection's weight.  K1 needs only w_att; when the forward head forks (g_head_overlap)
// the other 80 MB of packing traffic go out on the side stream, BESIDE the tensor-bound K1 instead of in front of it.
template <typename T>
int prep_weights(const sat_decoder_args& a, DecWs<T>& w, cudaStream_t st, cudaStream_t rest) {
    const int D = a.D, H = a.H, V = a.V;
    misc::PackList<T> pl;       // the packing jobs of a stream go out in one launch
    if (a.attention) {
        SAT_TRY(pl.weight(w.w_att, D, a.att_W_w, D, H, D, 0, 0));
        if (rest != st) SAT_TRY(pl.flush(st));
        SAT_TRY(pl.weight(w.w_uf, H, a.att_U_w, H, H, H, 0, 0));                       // rows [0,H): U
        SAT_TRY(pl.weight(w.w_uf + (size_t)H * H, H, a.f_beta_w, H, D, H, 0, 0));       // rows [H,H+D): f_beta
        SAT_TRY(pl.bias(w.b_uf, a.att_U_b, nullptr, H, 0));
        SAT_TRY(pl.bias(w.b_uf + H, a.f_beta_b, nullptr, D, 0));
    }
    // gate weights, rows interleaved (4j+g).  lstm.weight_ih columns: [0,H) embedding, [H,H+D) context.
    SAT_TRY(pl.weight(w.w_gx, H, a.lstm_w_ih, H + D, 4 * H, H, 0, H));
    SAT_TRY(pl.weight(w.w_gzh, D + H, a.lstm_w_ih, H + D, 4 * H, D, H, H));            // cols [0,D): context
    SAT_TRY(pl.weight(w.w_gzh + D, D + H, a.lstm_w_hh, H, 4 * H, H, 0, H));            // cols [D,D+H): hidden
    SAT_TRY(pl.bias(w.b_g, a.lstm_b_ih, a.lstm_b_hh, 4 * H, H));
    SAT_TRY(pl.weight(w.w_init, D, a.init_h_w, D, H, D, 0, 0));
    SAT_TRY(pl.weight(w.w_init + (size_t)H * D, D, a.init_c_w, D, H, D, 0, 0));
    SAT_TRY(pl.bias(w.b_init, a.init_h_b, nullptr, H, 0));
    SAT_TRY(pl.bias(w.b_init + H, a.init_c_b, nullptr, H, 0));
    if (a.ado) {
        SAT_REQUIRE(a.f_h_w && a.f_z_w && a.f_out_w, "ado=1 needs f_h/f_z/f_out parameters");
        SAT_TRY(pl.weight(w.w_fh, H, a.f_h_w, H, H, H, 0, 0));
        SAT_TRY(pl.weight(w.w_fz, D, a.f_z_w, D, H, D, 0, 0));
        SAT_TRY(pl.weight(w.w_out, H, a.f_out_w, H, V, H, 0, 0));
    } else {
        SAT_TRY(pl.weight(w.w_out, H, a.deep_w, H, V, H, 0, 0));
    }
    return pl.flush(rest);
}

// output layer for steps [t0, t0+nt): nt == T (teacher forcing: one GEMM over all (t,b) rows) or nt == 1.
template <typename T>
int output_layer(const sat_decoder_args& a, DecWs<T>& w, cudaStream_t st, int t0, int nt) {
    const int B = a.B, D = a.D, H = a.H, V = a.V, Tn = a.T;
    const int M = nt * B;
    float* logits = a.preds ? a.preds : w.logits_tmp;            // (B,T,ldl)
    const int64_t ldl = w.ldl;
    const T* hd = w.Hdrop + (size_t)t0 * B * H;
    auto set_out = [&](EpiLinear<T, IsF32<T>::value>& e) {
        if (nt == 1) { e.Cf = logits + (size_t)t0 * ldl; e.ldcf = (int64_t)Tn * ldl; }
        else { e.Cf = logits; e.ldcf = ldl; e.remapB = B; e.remapT = Tn; }
        e.vec_ok = e.vec_ok && (ldl % 8 == 0);
    };
    if (!a.ado) {
        auto e = lin<T>(M, V, a.deep_b, SAT_ACT_NONE, logits, ldl, nullptr, 0);
        set_out(e);
        return gemm<T>(st, hd, H, true, w.w_out, H, true, M, V, H, e);
    }
    // advanced deep output (decoder.py:149-158): relu(f_out(relu(f_h h) + relu(f_z ctx) + emb)), un-gated ctx (Q8)
    T* p1 = w.P1 + (size_t)t0 * B * H; T* p2 = w.P2 + (size_t)t0 * B * H; T* cb = w.Comb + (size_t)t0 * B * H;
    const T* ctx = w.Ctx + (size_t)t0 * B * D; const T* emb = w.E + (size_t)t0 * B * H;
    auto e1 = lin<T>(M, H, a.f_h_b, SAT_ACT_RELU, nullptr, 0, p1, H);
    SAT_TRY(gemm<T>(st, hd, H, true, w.w_fh, H, true, M, H, H, e1));
    auto e2 = lin<T>(M, H, a.f_z_b, SAT_ACT_RELU, nullptr, 0, p2, H);
    SAT_TRY(gemm<T>(st, ctx, D, true, w.w_fz, D, true, M, H, D, e2));
    misc::add3_kernel<T><<<misc::grid_for((int64_t)M * H), 256, 0, st>>>(cb, p1, p2, emb, (int64_t)M * H);
    SAT_CHECK_CUDA(cudaGetLastError());
    auto e3 = lin<T>(M, V, a.f_out_b, SAT_ACT_RELU, logits, ldl, nullptr, 0);
    set_out(e3);
    return gemm<T>(st, cb, H, true, w.w_out, H, true, M, V, H, e3);
}

// 2 when the recurrent loop runs as two half-batches on two streams (see SideStream above), else 1
template <typename T>
inline int dec_split(const sat_decoder_args& a, const DecWs<T>& w) {
    return (g_dec_split && a.tf && a.attention && w.attn_stream && a.B >= 128 && a.B % 16 == 0) ? 2 : 1;
}

template <typename T>
int decoder_forward_t(const sat_decoder_args& a) {
    constexpr bool PRECISE = IsF32<T>::value;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(a.stream);
    DecWs<T> w;
    const int64_t need = carve<T>(a, w, reinterpret_cast<uint8_t*>(a.workspace));
    SAT_REQUIRE(a.workspace && a.workspace_bytes >= need, "workspace too small: %lld < %lld", (long long)a.workspace_bytes, (long long)need);
    SAT_REQUIRE(((uintptr_t)a.workspace & 255) == 0, "workspace must be 256B aligned");
    const int B = a.B, L = a.L, D = a.D, H = a.H, V = a.V, Tn = a.T;
    const int64_t zw = D + H;
    SAT_CHECK_CUDA(cudaMemsetAsync(w.flags, 0, 4 * sizeof(int), st));
    // features in compute dtype, contiguous (Q20)
    if (reinterpret_cast<const void*>(w.a_c) != a.img_features) {
        if (a.a_dtype == SAT_DTYPE_F32)
            SAT_TRY((misc::cast3d<float, T>(st, w.a_c, reinterpret_cast<const float*>(a.img_features), B, L, D, a.a_stride_b, a.a_stride_l, a.a_stride_d)));
        else
            SAT_TRY((misc::cast3d<bf16, T>(st, w.a_c, reinterpret_cast<const bf16*>(a.img_features), B, L, D, a.a_stride_b, a.a_stride_l, a.a_stride_d)));
    }
    // Before the loop, two independent pieces: K1 (tensor-bound, 42 registers per thread, one CTA per SM) on the caller's stream
    // and, with attention on, everything else -- mean over L, initial state, embeddings and their gate share -- on the side
    // stream.  The HBM-bound mean_l blocks fit on the SMs BESIDE the K1 CTAs and read the same features K1 is pulling through L2.
    const bool head_overlap = g_head_overlap && a.attention;
    SideStream* hd = nullptr;
    cudaStream_t sh = st;
    if (head_overlap) {
        SAT_TRY(side_stream(&hd));
        sh = hd->s;
        SAT_CHECK_CUDA(cudaEventRecord(hd->fork, st));
        SAT_CHECK_CUDA(cudaStreamWaitEvent(sh, hd->fork, 0));
    }
    SAT_TRY(prep_weights<T>(a, w, st, g_pack_split ? sh : st));
    if (a.attention) {
        // K1: hoisted encoder projection Wa = a W^T + b_W (attention.py:16, Q3)
        auto e = lin<T>(B * L, H, a.att_W_b, SAT_ACT_NONE, nullptr, 0, w.wa, H);
        SAT_TRY(gemm<T>(st, w.a_c, D, true, w.w_att, D, true, B * L, H, D, e));
    }
    // initial state (decoder.py:137-147): [h0 | c0] = tanh(abar [init_h; init_c]^T + b)
    SAT_TRY(misc::mean_l<T>(sh, w.a_c, w.abar_f, w.abar, B, L, D, w.mean_part, (int64_t)4 * B * D));
    {
        auto e = lin<T>(B, 2 * H, w.b_init, SAT_ACT_TANH, w.hc0, 2 * H, nullptr, 0);
        SAT_TRY(gemm<T>(sh, w.abar, D, true, w.w_init, D, true, B, 2 * H, D, e));
        misc::init_state_scatter_kernel<T><<<sat_cdiv(B * H, 256), 256, 0, sh>>>(w.hc0, w.ZH + D, zw, w.Cst, H, B, H);
        SAT_CHECK_CUDA(cudaGetLastError());
    }
    if (!a.attention) {
        // decoder.py:101-105: alpha = 1/L, context = mean(a), no gate; identical for every step
        dim3 g(Tn, B);
        misc::bcast_rows_kernel<T><<<g, 128, 0, sh>>>(w.ZH, (int64_t)B * zw, zw, w.abar, D, Tn);
        misc::bcast_rows_kernel<T><<<g, 128, 0, sh>>>(w.Ctx, (int64_t)B * D, D, w.abar, D, Tn);
        misc::fill_kernel<<<misc::grid_for((int64_t)B * Tn * L), 256, 0, sh>>>(a.alphas, (int64_t)B * Tn * L, 1.0f / (float)L);
        SAT_CHECK_CUDA(cudaGetLastError());
    }
    // embeddings (decoder.py:85-89) and their share of the gate pre-activations, batched over t when teacher forcing
    if (a.tf) {
        misc::gather_bt_kernel<T><<<B * Tn, 128, 0, sh>>>(w.E, a.embedding, H, a.captions, a.cap_stride, B, V, w.flags);
        SAT_CHECK_CUDA(cudaGetLastError());
        auto e = lin<T>(B * Tn, 4 * H, w.b_g, SAT_ACT_NONE, w.Xg, 4 * H, nullptr, 0);
        SAT_TRY(gemm<T>(sh, w.E, H, true, w.w_gx, H, true, B * Tn, 4 * H, H, e));
    } else {
        misc::fill_tokens_kernel<<<sat_cdiv(B, 128), 128, 0, sh>>>(w.tok_in, B, (int64_t)a.start_token);
        SAT_CHECK_CUDA(cudaGetLastError());
    }
    if (head_overlap) {
        SAT_CHECK_CUDA(cudaEventRecord(hd->join, sh));
        SAT_CHECK_CUDA(cudaStreamWaitEvent(st, hd->join, 0));
    }
    // one decode step for the captions [b0, b0 + nb) on stream s
    auto fwd_step = [&](int t, int b0, int nb, cudaStream_t s, uint8_t* scratch, cudaEvent_t wait_ev, cudaEvent_t rec_ev) -> int {
        const int gsm = nb < B ? g_split_gemm_sms : 0, actas = nb < B ? g_split_attn_ctas : 0;
        T* zh_t = w.ZH + ((size_t)t * B + b0) * zw;
        if (a.attention) {
            float* uf_t = w.UF + ((size_t)t * B + b0) * (H + D);
            auto e = lin<T>(nb, H + D, w.b_uf, SAT_ACT_NONE, uf_t, H + D, nullptr, 0);
            SAT_TRY(gemm<T>(s, zh_t + D, zw, true, w.w_uf, H, true, nb, H + D, H, e, 0, gsm, 1));       // [U h | f_beta h]
            attn::FwdParams p;
            memset(&p, 0, sizeof(p));
            p.a = w.a_c + (size_t)b0 * L * D; p.wa = w.wa + (size_t)b0 * L * H;
            p.uf = uf_t; p.ld_uf = H + D; p.v = a.att_v_w; p.bv = a.att_v_b;
            p.alpha = a.alphas + ((size_t)b0 * Tn + t) * L; p.ld_alpha = (int64_t)Tn * L;
            p.ctx = w.Ctx + ((size_t)t * B + b0) * D; p.ld_ctx = D;
            p.z = zh_t; p.ld_z = zw;
            p.NI = nb; p.NB = 1; p.L = L; p.D = D; p.H = H; p.gate = a.gate; p.max_ctas = actas;
            if (w.attn_stream) {
                if (a.need_grad) { p.ctx32 = w.Ctx32 + ((size_t)t * B + b0) * D; p.ld_c32 = D; }
                p.static_inputs = 1;         // a_c / wa were written before the loop
                p.s_part = reinterpret_cast<float*>(scratch);
            }
            if (wait_ev) SAT_CHECK_CUDA(cudaStreamWaitEvent(s, wait_ev, 0));      // the other half's attention kernel has finished
            SAT_TRY((attn_step_fwd<T, PRECISE>(p, s)));
            if (rec_ev) SAT_CHECK_CUDA(cudaEventRecord(rec_ev, s));
        }
        // K3: gates = Xg[t] + [z | h] [W_ih[:,H:] | W_hh]^T, LSTM cell in the epilogue (decoder.py:108-115)
        EpiLstm<T, PRECISE> e;
        e.xg = w.Xg + ((size_t)t * B + b0) * 4 * H; e.ld_xg = 4 * H; e.xg_rows = nullptr; e.xg2 = nullptr; e.ld_xg2 = 0; e.bias = nullptr;
        e.c_prev = w.Cst + ((size_t)t * B + b0) * H; e.ld_cp = H;
        e.c_out = w.Cst + ((size_t)(t + 1) * B + b0) * H; e.ld_co = H;
        e.h_next = w.ZH + ((size_t)(t + 1) * B + b0) * zw + D; e.ld_hn = zw;
        e.h_out = w.Hdrop + ((size_t)t * B + b0) * H; e.ld_ho = H;
        e.mask = a.dropout_mask ? a.dropout_mask + ((size_t)b0 * Tn + t) * H : nullptr; e.ld_mask = (int64_t)Tn * H;
        e.gates = a.need_grad ? w.Gates + ((size_t)t * B + b0) * 4 * H : nullptr; e.ld_g = 4 * H;
        e.M = nb; e.H = H;
        return gemm<T>(s, zh_t, zw, true, w.w_gzh, zw, true, nb, 4 * H, D + H, e, 0, gsm, 1);
    };
    const int nsplit = dec_split(a, w);
    SideStream* ss = nullptr;
    if (nsplit > 1) {
        SAT_TRY(side_stream(&ss));
        SAT_CHECK_CUDA(cudaEventRecord(ss->fork, st));
        SAT_CHECK_CUDA(cudaStreamWaitEvent(ss->s, ss->fork, 0));
    }
    for (int t = 0; t < Tn; ++t) {
        if (!a.tf) {
            // token fed at step t: start token, then argmax of the previous logits (decoder.py:79-82,132-133)
            SAT_TRY(misc::gather_rows<T>(st, w.E + (size_t)t * B * H, H, a.embedding, H, w.tok_in + (size_t)t * B, 1, B, V));
            auto e = lin<T>(B, 4 * H, w.b_g, SAT_ACT_NONE, w.Xg + (size_t)t * B * 4 * H, 4 * H, nullptr, 0);
            SAT_TRY(gemm<T>(st, w.E + (size_t)t * B * H, H, true, w.w_gx, H, true, B, 4 * H, H, e));
        }
        if (nsplit > 1) {
            const int nb0 = B / 2;
            // chained: the two halves' attention kernels take turns on the HBM (A_t, B_t, A_t+1, ...) instead of drifting
            // into lockstep, so each half's GEMMs run while the other half streams
            const bool ch = g_split_chain != 0;
            SAT_TRY(fwd_step(t, 0, nb0, st, w.attn_scratch, (ch && t > 0) ? ss->attn_done[1] : nullptr, ch ? ss->attn_done[0] : nullptr));
            SAT_TRY(fwd_step(t, nb0, B - nb0, ss->s, w.attn_scratch2, ch ? ss->attn_done[0] : nullptr, ch ? ss->attn_done[1] : nullptr));
        } else {
            SAT_TRY(fwd_step(t, 0, B, st, w.attn_scratch, nullptr, nullptr));
        }
        if (!a.tf) {
            SAT_TRY(output_layer<T>(a, w, st, t, 1));
            if (t + 1 < Tn) {
                float* logits = (a.preds ? a.preds : w.logits_tmp) + (size_t)t * w.ldl;
                misc::argmax_rows_kernel<<<B, 256, 0, st>>>(logits, (int64_t)Tn * w.ldl, V, w.tok_in + (size_t)(t + 1) * B, 1);
                SAT_CHECK_CUDA(cudaGetLastError());
            }
        }
    }
    if (nsplit > 1) {
        SAT_CHECK_CUDA(cudaEventRecord(ss->join, ss->s));
        SAT_CHECK_CUDA(cudaStreamWaitEvent(st, ss->join, 0));
    }
    if (a.tf) SAT_TRY(output_layer<T>(a, w, st, 0, Tn));   // K4: one (T*B) x H x V GEMM after the loop
    if (a.fused_loss) {
        SAT_REQUIRE(a.loss != nullptr, "fused_loss needs a loss output");
        const float* logits = a.preds ? a.preds : w.logits_tmp;
        // with need_grad the same pass also leaves d loss / d logits for a unit upstream gradient (what the fused train step
        // feeds back); sat_decoder_backward recomputes it only when it is handed a dloss scalar
        misc::ce_rows<T>(st, B * Tn, logits, w.ldl, V, Tn, a.captions, a.cap_stride, w.rowloss,
                         a.need_grad ? w.dlog : (T*)nullptr, w.Vp, nullptr, 1.0f / (float)(B * (Tn - 1)), a.ado, w.flags);
        misc::loss_partial_kernel<<<misc::LOSS_PARTS, 256, 0, st>>>(w.rowloss, B * Tn, a.alphas, B, Tn, L, w.loss_part);
        misc::loss_finalize_kernel<<<1, 256, 0, st>>>(w.loss_part, misc::LOSS_PARTS, B * (Tn - 1), a.alphas != nullptr, B, L, a.alpha_c, a.loss,
                                                      w.flags);
        SAT_CHECK_CUDA(cudaGetLastError());
    } else if (a.loss) {
        misc::status_kernel<<<1, 32, 0, st>>>(a.loss, w.flags);
        SAT_CHECK_CUDA(cudaGetLastError());
    }
    return SAT_OK;
}

// ---------------------------------------------------------------------------------------------------
// backward (BPTT)
// ---------------------------------------------------------------------------------------------------
template <typename T>
int decoder_backward_t(const sat_decoder_args& a, const sat_decoder_grads& g) {
    constexpr bool PRECISE = IsF32<T>::value;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(a.stream);
    SAT_REQUIRE(a.need_grad, "forward was run without need_grad");
    DecWs<T> w;
    const int64_t need = carve<T>(a, w, reinterpret_cast<uint8_t*>(a.workspace));
    SAT_REQUIRE(a.workspace && a.workspace_bytes >= need, "workspace too small");
    const int B = a.B, L = a.L, D = a.D, H = a.H, V = a.V, Tn = a.T;
    const int64_t zw = D + H;
    const int BT = B * Tn;
    const int64_t Vp = w.Vp;
    const float* logits = a.preds ? a.preds : w.logits_tmp;

    // ---- seed: dlogits (T dtype, time-major rows, row pitch Vp) ------------------------------------
    const float* galpha = g.dalphas;
    if (a.fused_loss) {
        const float scale = 1.0f / (float)(B * (Tn - 1));
        if (g.dloss)     // otherwise the forward's unit-gradient dlogits are still in the workspace
            misc::ce_rows<T>(st, B * Tn, logits, w.ldl, V, Tn, a.captions, a.cap_stride, nullptr, w.dlog, Vp, g.dloss, scale, a.ado, nullptr);
        if (a.attention) {
            misc::alpha_reg_grad_kernel<<<sat_cdiv(B * L, 128), 128, 0, st>>>(a.alphas, w.galpha, B, Tn, L, a.alpha_c, g.dloss);
            galpha = w.galpha;
        }
        SAT_CHECK_CUDA(cudaGetLastError());
    } else {
        SAT_REQUIRE(g.dpreds != nullptr, "dpreds is NULL");
        misc::dpreds_to_dlog_kernel<T><<<BT, 256, 0, st>>>(w.dlog, Vp, g.dpreds, a.ado ? logits : nullptr, w.ldl, B, Tn, V);
        SAT_CHECK_CUDA(cudaGetLastError());
    }

    // ---- output layer backward, batched over all (t,b) ---------------------------------------------
    float* g_out_w = a.ado ? g.f_out_w : g.deep_w;
    float* g_out_b = a.ado ? g.f_out_b : g.deep_b;
    const T* out_in = a.ado ? w.Comb : w.Hdrop;     // input rows of the V-projection
    // (dW_out = dlog^T X and db_out are not needed by the BPTT loop: they run in the side chain of the tail below)
    if (!a.ado) {
        auto e = lin<T>(BT, H, nullptr, SAT_ACT_NONE, w.dHd, H, nullptr, 0);
        SAT_TRY(gemm<T>(st, w.dlog, Vp, true, w.w_out, H, false, BT, H, V, e));          // dHdrop = dlog W_out
    } else {
        // dcomb = dlog W_fout ; dP1 = dcomb*relu'(P1) ; dP2 = dcomb*relu'(P2) ; dE += dcomb
        auto e = lin<T>(BT, H, nullptr, SAT_ACT_NONE, w.dcomb, H, nullptr, 0);
        SAT_TRY(gemm<T>(st, w.dlog, Vp, true, w.w_out, H, false, BT, H, V, e));
        misc::relu_bwd_kernel<T><<<misc::grid_for((int64_t)BT * H), 256, 0, st>>>(w.dP1, w.dcomb, w.P1, (int64_t)BT * H);
        misc::relu_bwd_kernel<T><<<misc::grid_for((int64_t)BT * H), 256, 0, st>>>(w.dP2, w.dcomb, w.P2, (int64_t)BT * H);
        SAT_CHECK_CUDA(cudaGetLastError());
        if (g.f_h_w) { auto e1 = lin<T>(H, H, nullptr, SAT_ACT_NONE, g.f_h_w, H, nullptr, 0);
            SAT_TRY(gemm<T>(st, w.dP1, H, false, w.Hdrop, H, false, H, H, BT, e1)); }
        if (g.f_h_b) SAT_TRY(misc::colsum<T>(st, w.dP1, H, BT, H, g.f_h_b, w.csum, w.csum_floats));
        if (g.f_z_w) { auto e1 = lin<T>(H, D, nullptr, SAT_ACT_NONE, g.f_z_w, D, nullptr, 0);
            SAT_TRY(gemm<T>(st, w.dP2, H, false, w.Ctx, D, false, H, D, BT, e1)); }
        if (g.f_z_b) SAT_TRY(misc::colsum<T>(st, w.dP2, H, BT, H, g.f_z_b, w.csum, w.csum_floats));
        auto e2 = lin<T>(BT, H, nullptr, SAT_ACT_NONE, w.dHd, H, nullptr, 0);
        SAT_TRY(gemm<T>(st, w.dP1, H, true, w.w_fh, H, false, BT, H, H, e2));            // dHdrop = dP1 W_fh
        auto e3 = lin<T>(BT, D, nullptr, SAT_ACT_NONE, w.dctx_all, D, nullptr, 0);
        SAT_TRY(gemm<T>(st, w.dP2, H, true, w.w_fz, D, false, BT, D, H, e3));            // dctx (un-gated) = dP2 W_fz
    }

    // ---- BPTT loop ------------------------------------------------------------------------------------
    SAT_CHECK_CUDA(cudaMemsetAsync(w.dh_carry, 0, (size_t)B * H * 4, st));
    SAT_CHECK_CUDA(cudaMemsetAsync(w.dc_carry, 0, (size_t)B * H * 4, st));
    // one BPTT step for the captions [b0, b0 + nb) on stream s
    auto bwd_step = [&](int t, int b0, int nb, cudaStream_t s, uint8_t* scratch, cudaEvent_t wait_ev, cudaEvent_t rec_ev) -> int {
        const int gsm = nb < B ? g_split_gemm_sms : 0, actas = nb < B ? g_split_attn_ctas : 0;
        const size_t r = (size_t)t * B + b0;       // first time-major row of this half at step t
        // need_dfeat keeps every step's d[z | h] (the post-loop d img_features pass re-reads the dctx columns)
        float* dzh = w.dZH + ((a.need_dfeat ? (size_t)t * B : (size_t)0) + b0) * zw;
        float* dh_carry = w.dh_carry + (size_t)b0 * H;
        // The cell backward of step t runs in the epilogue of the previous iteration's dh GEMM (EpiDhCell) when attention is on;
        // only the first iteration (t = T-1, dh_carry = 0) and the no-attention decoder launch the pointwise kernel.
        const bool cell_fused = g_fuse_cell_bwd && a.attention;
        if (!cell_fused || t == Tn - 1) {
            cudaLaunchConfig_t cfg;
            cudaLaunchAttribute attr[1];
            pdl_config(&cfg, attr, dim3(sat_cdiv(H, 128), nb), dim3(128), 0, s);
            SAT_CHECK_CUDA(cudaLaunchKernelEx(
                &cfg, misc::lstm_bwd_kernel<T>, (const T*)(w.Gates + r * 4 * H), (int64_t)(4 * H),
                (const float*)(w.Cst + r * H), (int64_t)H, (const float*)(w.Cst + ((size_t)(t + 1) * B + b0) * H), (int64_t)H,
                (const float*)(w.dHd + r * H), (int64_t)H,
                (const float*)(a.dropout_mask ? a.dropout_mask + ((size_t)b0 * Tn + t) * H : nullptr), (int64_t)Tn * H,
                (const float*)dh_carry, w.dc_carry + (size_t)b0 * H, w.dG + r * 4 * H, (int64_t)(4 * H), nb, H));
        }
        // d[z | h_prev] = dG [W_ih[:,H:] | W_hh]   (streaming attention: with the context gate's backward in the epilogue)
        if (a.attention && w.attn_stream) {
            EpiDzGate<T, PRECISE> e;
            e.out = dzh; e.ld_out = zw;
            e.uf = w.UF + r * (H + D); e.ld_uf = H + D;
            e.ctx32 = w.Ctx32 + r * D; e.ld_c32 = D;
            e.extra = a.ado ? w.dctx_all + r * D : nullptr; e.ld_ex = D;
            e.duf = w.dUF + r * (H + D); e.ld_duf = H + D;
            e.M = nb; e.D = D; e.H = H; e.gate = a.gate;
            SAT_TRY(gemm<T>(s, w.dG + r * 4 * H, 4 * H, true, w.w_gzh, zw, false, nb, D + H, 4 * H, e, 0, gsm, 1));
        } else {
            auto e = lin<T>(nb, D + H, nullptr, SAT_ACT_NONE, dzh, zw, nullptr, 0);
            SAT_TRY(gemm<T>(s, w.dG + r * 4 * H, 4 * H, true, w.w_gzh, zw, false, nb, D + H, 4 * H, e, 0, gsm, 1));
        }
        if (a.attention) {
            attn::BwdParams p;
            memset(&p, 0, sizeof(p));
            p.a = w.a_c + (size_t)b0 * L * D; p.wa = w.wa + (size_t)b0 * L * H; p.uf = w.UF + r * (H + D); p.ld_uf = H + D;
            p.v = a.att_v_w; p.alpha = a.alphas + ((size_t)b0 * Tn + t) * L; p.ld_alpha = (int64_t)Tn * L;
            p.ctx = w.Ctx + r * D; p.ld_ctx = D;
            p.dz = dzh; p.ld_dz = zw;
            p.dctx_extra = a.ado ? w.dctx_all + r * D : nullptr; p.ld_dce = D;
            p.galpha = galpha ? galpha + ((size_t)b0 * Tn + t) * L : nullptr; p.ld_ga = (int64_t)Tn * L;
            p.duf = w.dUF + r * (H + D); p.ld_duf = H + D;
            p.dwa_acc = w.dwa_acc ? w.dwa_acc + (size_t)b0 * L * H : nullptr;
            p.dv_acc = w.dv_acc + (size_t)b0 * H; p.dbv_acc = w.dbv_acc + b0;
            // generic kernels: sum_t alpha_t (x) dctx_t accumulates straight into the caller's gradient buffer; the streaming
            // kernels leave that term to dfeat_kernel after the loop (dctx_t stays in dZH[t])
            p.da_acc = (g.img_features && !w.attn_stream) ? g.img_features + (size_t)b0 * L * D : nullptr;
            if (w.attn_stream) {
                if (w.de_all) p.de_out = w.de_all + ((size_t)t * B + b0) * L;          // sum_t dpre is rebuilt after the loop
                else p.dpre_out = w.dpre_all + ((size_t)t * B + b0) * L * H;
                p.ctx32 = w.Ctx32 + r * D; p.ld_c32 = D;
                p.s_part = reinterpret_cast<float*>(scratch);
                p.dz_is_dctx = 1;            // EpiDzGate above already turned dz into dctx and wrote d f_beta
                p.static_inputs = 1;
            }
            p.first = (t == Tn - 1);
            p.B = nb; p.L = L; p.D = D; p.H = H; p.gate = a.gate; p.max_ctas = actas;
            if (wait_ev) SAT_CHECK_CUDA(cudaStreamWaitEvent(s, wait_ev, 0));
            SAT_TRY((attn_step_bwd<T, PRECISE>(p, s)));
            if (rec_ev) SAT_CHECK_CUDA(cudaEventRecord(rec_ev, s));
            // dh_prev = dZH[:, D:] + [dUh | dfb] [U ; f_beta]
            if (cell_fused && t > 0) {
                const size_t rp = (size_t)(t - 1) * B + b0;      // rows of step t-1
                EpiDhCell<T, PRECISE> e;
                e.addend = dzh + D; e.ld_add = zw;
                e.gates = w.Gates + rp * 4 * H; e.ld_g = 4 * H;
                e.c_prev = w.Cst + rp * H; e.ld_cp = H;
                e.c_new = w.Cst + r * H; e.ld_cn = H;            // Cst[t] = state leaving step t-1
                e.dhd = w.dHd + rp * H; e.ld_dhd = H;
                e.mask = a.dropout_mask ? a.dropout_mask + ((size_t)b0 * Tn + (t - 1)) * H : nullptr; e.ld_mask = (int64_t)Tn * H;
                e.dc_carry = w.dc_carry + (size_t)b0 * H; e.ld_dc = H;
                e.dG = w.dG + rp * 4 * H; e.ld_dg = 4 * H;
                e.M = nb; e.H = H;
                SAT_TRY(gemm<T>(s, w.dUF + r * (H + D), H + D, true, w.w_uf, H, false, nb, H, H + D, e, 0, gsm, 1));
            } else {
                auto e = lin<T>(nb, H, nullptr, SAT_ACT_NONE, dh_carry, H, nullptr, 0, dzh + D, zw);
                SAT_TRY(gemm<T>(s, w.dUF + r * (H + D), H + D, true, w.w_uf, H, false, nb, H, H + D, e, 0, gsm, 1));
            }
        } else {
            // no attention: z is the constant mean feature; only the hidden path carries on
            SAT_CHECK_CUDA(cudaMemcpy2DAsync(dh_carry, (size_t)H * 4, dzh + D, (size_t)zw * 4, (size_t)H * 4, nb,
                                             cudaMemcpyDeviceToDevice, s));
        }
        return SAT_OK;
    };
    const int nsplit = dec_split(a, w);
    SideStream* ss = nullptr;
    if (nsplit > 1) {
        SAT_TRY(side_stream(&ss));
        SAT_CHECK_CUDA(cudaEventRecord(ss->fork, st));
        SAT_CHECK_CUDA(cudaStreamWaitEvent(ss->s, ss->fork, 0));
    }
    for (int t = Tn - 1; t >= 0; --t) {
        if (nsplit > 1) {
            const int nb0 = B / 2;
            const bool ch = g_split_chain != 0;
            SAT_TRY(bwd_step(t, 0, nb0, st, w.attn_scratch, (ch && t < Tn - 1) ? ss->attn_done[1] : nullptr, ch ? ss->attn_done[0] : nullptr));
            SAT_TRY(bwd_step(t, nb0, B - nb0, ss->s, w.attn_scratch2, ch ? ss->attn_done[0] : nullptr, ch ? ss->attn_done[1] : nullptr));
        } else {
            SAT_TRY(bwd_step(t, 0, B, st, w.attn_scratch, nullptr, nullptr));
        }
    }
    if (nsplit > 1) {
        SAT_CHECK_CUDA(cudaEventRecord(ss->join, ss->s));
        SAT_CHECK_CUDA(cudaStreamWaitEvent(st, ss->join, 0));
    }

    // ---- after the loop: two independent chains on two streams ---------------------------------------------------
    // main chain (HBM / tensor bound): sum over t of the attention pre-activation gradients -> dW_att GEMM -> dv, db_v.
    // side chain (many small launches): initial-state, gate, embedding, [U; f_beta] and output-layer weight gradients.
    // They share no buffer (gscratch / csum2 belong to the side chain, csum to the main chain) and join before returning.
    SideStream* tl = nullptr;
    SAT_TRY(side_stream(&tl));
    cudaStream_t sb = g_tail_overlap ? tl->s : st;
    float* csb = g_tail_overlap ? w.csum2 : w.csum;
    // g_tail_overlap == 3: the side chain itself forks.  Its launches are small and mutually independent (each GEMM covers 64-80
    // of the 148 SMs), so embedding / [U; f_beta] / output-layer gradients run on a second side stream beside the initial-state
    // and gate-weight gradients, with their own scratch (gscratch2 / csum3).
    const bool three = g_tail_overlap == 3;
    cudaStream_t sc = three ? tl->s2 : sb;
    float* csc = three ? w.csum3 : csb;
    float* gsc = three ? w.gscratch2 : w.gscratch;
    if (g_tail_overlap) {
        SAT_CHECK_CUDA(cudaEventRecord(tl->fork, st));
        SAT_CHECK_CUDA(cudaStreamWaitEvent(sb, tl->fork, 0));
        if (three) SAT_CHECK_CUDA(cudaStreamWaitEvent(sc, tl->fork, 0));
    }
    cudaStream_t caller = st;
    const bool main_hp = g_tail_hp && g_tail_overlap && a.attention;
    if (main_hp) {          // the main chain moves to the high-priority stream (forked from the same event as the side chains)
        SAT_CHECK_CUDA(cudaStreamWaitEvent(tl->hp, tl->fork, 0));
        st = tl->hp;
    }
    if (a.attention) {      // main chain
        // encoder projection: dW (H x D) = (sum_t dpre)^T a ; db_W = colsum ; dv, db_v
        bool att_W_b_done = false;
        if (w.attn_stream) {
            if (w.de_all) {      // one pass over Wa: sum_t dpre_t rebuilt from de_t / Uh_t (+ the bias gradient's column sums)
                SAT_TRY((misc::dpre_sum<T, PRECISE>(st, w.dwa_t, w.wa, w.UF, H + D, w.de_all, a.att_v_w, B, L, H, Tn, g.att_W_b,
                                                    w.csum, w.csum_floats)));
                att_W_b_done = true;
            } else if (g.att_W_b) {     // one pass: sum over t + the bias gradient's column sums
                SAT_TRY(misc::sum_over_t_colsum<T>(st, w.dwa_t, w.dpre_all, (int64_t)B * L, H, Tn, g.att_W_b, w.csum, w.csum_floats, g_tail_overlap == 2));
                att_W_b_done = true;
            } else {
                SAT_TRY(misc::sum_over_t<T>(st, w.dwa_t, w.dpre_all, (int64_t)B * L * H, Tn));
            }
        } else {
            SAT_TRY((misc::cast<float, T>(st, w.dwa_t, w.dwa_acc, (int64_t)B * L * H)));
        }
        if (g.att_W_w) {
            auto e2 = lin<T>(H, D, nullptr, SAT_ACT_NONE, g.att_W_w, D, nullptr, 0);
            SAT_TRY(gemm<T>(st, w.dwa_t, H, false, w.a_c, D, false, H, D, B * L, e2));
        }
        if (g.att_W_b && !att_W_b_done) {
            if (w.attn_stream) SAT_TRY(misc::colsum<T>(st, w.dwa_t, H, B * L, H, g.att_W_b, w.csum, w.csum_floats));
            else SAT_TRY(misc::colsum<float>(st, w.dwa_acc, H, B * L, H, g.att_W_b, w.csum, w.csum_floats));
        }
        if (g.att_v_w) SAT_TRY(misc::colsum<float>(st, w.dv_acc, H, B, H, g.att_v_w, w.csum, w.csum_floats));
        if (g.att_v_b) SAT_TRY(misc::colsum<float>(st, w.dbv_acc, 1, B, 1, g.att_v_b, w.csum, w.csum_floats));
        if (g.img_features) {
            // d img_features, projection share: (sum_t dpre) W   ((B L) x D; added to the generic kernels' running sum)
            auto e3 = lin<T>(B * L, D, nullptr, SAT_ACT_NONE, g.img_features, D, nullptr, 0,
                             w.attn_stream ? nullptr : g.img_features, D);
            SAT_TRY(gemm<T>(st, w.dwa_t, H, true, w.w_att, D, false, B * L, D, H, e3));
        }
    }

    if (main_hp) {
        SAT_CHECK_CUDA(cudaEventRecord(tl->joinhp, st));
        st = caller;
        SAT_CHECK_CUDA(cudaStreamWaitEvent(st, tl->joinhp, 0));
    }
    // ---- side chain: initial state (decoder.py:137-147) ------------------------------------------------------------
    misc::dinit_kernel<T><<<sat_cdiv(B * H, 256), 256, 0, sb>>>(w.dpre0, w.dh_carry, w.dc_carry, w.hc0, B, H);
    SAT_CHECK_CUDA(cudaGetLastError());
    {
        auto e = lin<T>(2 * H, D, nullptr, SAT_ACT_NONE, w.gscratch, D, nullptr, 0);
        SAT_TRY(gemm<T>(sb, w.dpre0, 2 * H, false, w.abar, D, false, 2 * H, D, B, e));
        if (g.init_h_w) SAT_CHECK_CUDA(cudaMemcpyAsync(g.init_h_w, w.gscratch, (size_t)H * D * 4, cudaMemcpyDeviceToDevice, sb));
        if (g.init_c_w) SAT_CHECK_CUDA(cudaMemcpyAsync(g.init_c_w, w.gscratch + (size_t)H * D, (size_t)H * D * 4, cudaMemcpyDeviceToDevice, sb));
        SAT_TRY(misc::colsum<T>(sb, w.dpre0, 2 * H, B, 2 * H, w.gscratch, csb, w.csum_floats));
        if (g.init_h_b) SAT_CHECK_CUDA(cudaMemcpyAsync(g.init_h_b, w.gscratch, (size_t)H * 4, cudaMemcpyDeviceToDevice, sb));
        if (g.init_c_b) SAT_CHECK_CUDA(cudaMemcpyAsync(g.init_c_b, w.gscratch + H, (size_t)H * 4, cudaMemcpyDeviceToDevice, sb));
        if (g.img_features) {     // d abar = dpre0 [init_h ; init_c]   (decoder.py:139-145)
            auto e2 = lin<T>(B, D, nullptr, SAT_ACT_NONE, w.dabar, D, nullptr, 0);
            SAT_TRY(gemm<T>(sb, w.dpre0, 2 * H, true, w.w_init, D, false, B, D, 2 * H, e2));
        }
    }
    // ---- side chain: batched weight gradients (contraction over k = t*B + b) -----------------------------------------
    {
        // gate weights: dW_gzh (4H x (D+H)) = dG^T [Z | Hprev] ; dW_gx (4H x H) = dG^T E ; un-interleave rows
        auto e = lin<T>(4 * H, D + H, nullptr, SAT_ACT_NONE, w.gscratch, zw, nullptr, 0);
        SAT_TRY(gemm<T>(sb, w.dG, 4 * H, false, w.ZH, zw, false, 4 * H, D + H, BT, e));
        if (g.lstm_w_ih) SAT_TRY(misc::unpack_grad(sb, g.lstm_w_ih, H + D, w.gscratch, zw, 4 * H, D, H, H));
        if (g.lstm_w_hh) SAT_TRY(misc::unpack_grad(sb, g.lstm_w_hh, H, w.gscratch + D, zw, 4 * H, H, 0, H));
        auto e2 = lin<T>(4 * H, H, nullptr, SAT_ACT_NONE, w.gscratch, H, nullptr, 0);
        SAT_TRY(gemm<T>(sb, w.dG, 4 * H, false, w.E, H, false, 4 * H, H, BT, e2));
        if (g.lstm_w_ih) SAT_TRY(misc::unpack_grad(sb, g.lstm_w_ih, H + D, w.gscratch, H, 4 * H, H, 0, H));
        // The initial-state and LSTM weight gradients (62 % of the gradient bytes at c2) are final here, ~160 us before the
        // tail ends: data-parallel callers start their all-reduce on this event (sat_wait_early_grads) under the rest of it.
        SAT_CHECK_CUDA(cudaEventRecord(tl->early, sb));
        // d b_ih = d b_hh = column sums of dG: computed once, copied
        float* gb = g.lstm_b_ih ? g.lstm_b_ih : g.lstm_b_hh;
        if (gb) SAT_TRY(misc::colsum<T>(sb, w.dG, 4 * H, BT, 4 * H, gb, csb, w.csum_floats, H));
        if (g.lstm_b_ih && g.lstm_b_hh)
            SAT_CHECK_CUDA(cudaMemcpyAsync(g.lstm_b_hh, g.lstm_b_ih, (size_t)4 * H * 4, cudaMemcpyDeviceToDevice, sb));
    }
    // embedding: dE = dG W_gx (+ dcomb for ado), scatter-add by token (skipped when frozen / NULL: Q23)
    if (g.embedding) {
        SAT_CHECK_CUDA(cudaMemsetAsync(g.embedding, 0, (size_t)V * H * 4, sc));
        auto e = lin<T>(BT, H, nullptr, SAT_ACT_NONE, nullptr, 0, w.dE, H, a.ado ? w.dcomb : nullptr, H);
        SAT_TRY(gemm<T>(sc, w.dG, 4 * H, true, w.w_gx, H, false, BT, H, 4 * H, e));
        if (a.tf) misc::scatter_add_bt_kernel<T><<<BT, 128, 0, sc>>>(g.embedding, H, w.dE, a.captions, a.cap_stride, B, V);
        else misc::scatter_add_rows_kernel<T><<<BT, 128, 0, sc>>>(g.embedding, H, w.dE, w.tok_in, V);
        SAT_CHECK_CUDA(cudaGetLastError());
    }
    if (a.attention) {
        // [U ; f_beta] grads: dW_uf ((H+D) x H) = dUF^T Hprev  (Hprev = ZH[:, D:], steps 0..T-1)
        auto e = lin<T>(H + D, H, nullptr, SAT_ACT_NONE, gsc, H, nullptr, 0);
        SAT_TRY(gemm<T>(sc, w.dUF, H + D, false, w.ZH + D, zw, false, H + D, H, BT, e));
        if (g.att_U_w) SAT_CHECK_CUDA(cudaMemcpyAsync(g.att_U_w, gsc, (size_t)H * H * 4, cudaMemcpyDeviceToDevice, sc));
        if (g.f_beta_w) SAT_CHECK_CUDA(cudaMemcpyAsync(g.f_beta_w, gsc + (size_t)H * H, (size_t)D * H * 4, cudaMemcpyDeviceToDevice, sc));
        SAT_TRY(misc::colsum<T>(sc, w.dUF, H + D, BT, H + D, gsc, csc, w.csum_floats));
        if (g.att_U_b) SAT_CHECK_CUDA(cudaMemcpyAsync(g.att_U_b, gsc, (size_t)H * 4, cudaMemcpyDeviceToDevice, sc));
        if (g.f_beta_b) SAT_CHECK_CUDA(cudaMemcpyAsync(g.f_beta_b, gsc + H, (size_t)D * 4, cudaMemcpyDeviceToDevice, sc));
    }
    // output layer weights (deferred from before the loop)
    if (g_out_w) {
        auto e = lin<T>(V, H, nullptr, SAT_ACT_NONE, g_out_w, H, nullptr, 0);
        SAT_TRY(gemm<T>(sc, w.dlog, Vp, false, out_in, H, false, V, H, BT, e));          // dW = dlog^T X
    }
    if (g_out_b) SAT_TRY(misc::colsum<T>(sc, w.dlog, Vp, BT, V, g_out_b, csc, w.csum_floats));
    if (three) {
        SAT_CHECK_CUDA(cudaEventRecord(tl->join2, sc));
        SAT_CHECK_CUDA(cudaStreamWaitEvent(sb, tl->join2, 0));
    }
    // every gradient except the encoder projection's (attention.W.*, attention.v.*: main chain) is final here: data-parallel
    // callers all-reduce them, and the optimizer updates them, under the rest of the main chain (sat_wait_side_grads)
    SAT_CHECK_CUDA(cudaEventRecord(tl->side, sb));
    if (g_tail_overlap) {
        SAT_CHECK_CUDA(cudaEventRecord(tl->join, sb));
        SAT_CHECK_CUDA(cudaStreamWaitEvent(st, tl->join, 0));
    }
    if (g.img_features) {
        // d img_features (train.py:132 leaves the encoder outside no_grad): the remaining shares, after both chains.
        //   attention on : += sum_t alpha[b,t,l] dctx_t[b,d] (streaming kernels only)  + d abar[b,d] / L
        //   attention off: z_t = ctx_t = abar for every t (decoder.py:101-105): (d abar + sum_t (dz_t + dctx_t)) / L
        SAT_REQUIRE(a.need_dfeat, "grads->img_features needs args->need_dfeat = 1 in both calls");
        misc::DfeatParams q;
        q.da = g.img_features; q.alphas = a.alphas; q.dzh = w.dZH; q.ld_dzh = zw; q.dctx_extra = a.ado ? w.dctx_all : nullptr;
        q.dabar = w.dabar; q.B = B; q.Tn = Tn; q.L = L; q.D = D;
        q.mode = !a.attention ? 2 : (w.attn_stream ? 1 : 0);
        dim3 grid(sat_cdiv(D, 256), B);
        misc::dfeat_kernel<<<grid, 256, 0, st>>>(q);
        SAT_CHECK_CUDA(cudaGetLastError());
    }
    return SAT_OK;
}
