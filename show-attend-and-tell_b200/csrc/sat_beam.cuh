// Batched beam search: Decoder.caption (decoder.py:160-269) for NI images at once, each with its own
// shrinking beam (Q16), raw-logit cumulative scores (Q13), step-1 row-0 rule (Q14), the reference's stop
// tokens (Q15), 51-iteration cap (Q17) and the "no completed sentence" fallback (Q18).
// Rows r = img*beam + slot share the image's staged features in the attention kernel (NB = beam).
// All bookkeeping (top-k over beam x V, completion, parent gather, back-pointers) stays on the device.
#pragma once
#include "sat_decoder.cuh"

namespace beam {

// K5: per image, the `n` best of (live rows x V) cumulative scores, sorted (score desc, flat index asc),
// followed by the reference's bookkeeping (decoder.py:206-250).
struct StepParams {
    const float* logits;     // (R, ldl): row pitch padded to a multiple of 8 floats so the GEMM epilogue stores vectors
    int64_t ldl;
    float* top_preds;        // (R)  cumulative score per live slot (in/out)
    int64_t* prev_word;      // (R)  word fed to the next step (out)
    int* live;               // (NI) current beam size (in/out)
    int* done;               // (NI)
    int* n_active;           // (1) images still running
    int* parent;             // (R)  for the state gather: slot j of the next step continues row parent[j] (out)
    int* cand_word;          // (S, NI, beam)
    int* cand_parent;        // (S, NI, beam)
    int* slot_cand;          // (S, NI, beam)
    float* comp_score;       // (NI, beam)
    int* comp_step;          // (NI, beam)
    int* comp_cand;          // (NI, beam)
    // state gather fused into this kernel (decoder.py:246-247): h / c rows of the surviving beams for the next step's operands
    void* X;                 // (R, ldx) T: next step's GEMM operand rows; the h part starts at column hoff
    int64_t ldx;
    int hoff;
    const void* Hn;          // (R, H) T: h of this step
    float* Cc;               // (R, H) cell state of the next step (out)
    const float* Cn;         // (R, H) cell state of this step
    int H, esize;            // hidden size, sizeof(T)
    int* ncomp;              // (NI)
    int* last_step;          // (NI) last executed step
    int* last_n;             // (NI) rows alive at the start of the last executed step
    int NI, beam, V, step, max_steps, stop_a, stop_b;
    int cache_floats;        // dynamic shared memory available for the score cache (floats)
};

constexpr int TOPK_THREADS = 512;
constexpr int MAXBEAM = 16;

__global__ void __launch_bounds__(TOPK_THREADS)
topk_step_kernel(StepParams p) {
    __shared__ float s_val[TOPK_THREADS / 32];
    __shared__ int s_idx[TOPK_THREADS / 32];
    __shared__ float pick_val[MAXBEAM];
    __shared__ int pick_idx[MAXBEAM];
    __shared__ float s_tp[MAXBEAM];
    __shared__ int s_parent[MAXBEAM];
    __shared__ int s_nl;
    extern __shared__ float s_cache[];      // cumulative scores of this image, filled by the first selection pass
    const int img = blockIdx.x;
    if (p.done[img]) return;
    const int n = p.live[img];
    const int rows = (p.step == 1) ? 1 : n;                    // decoder.py:206-209 (Q14)
    const int total = rows * p.V;
    const bool cached = total <= p.cache_floats;   // later passes then read shared memory instead of global
    const float* lg = p.logits + (size_t)img * p.beam * p.ldl;
    if (threadIdx.x < MAXBEAM) s_tp[threadIdx.x] = (threadIdx.x < n) ? p.top_preds[img * p.beam + threadIdx.x] : 0.0f;
    __syncthreads();
    float last_v = INFINITY;
    int last_i = -1;
    if (cached) {
        // Load phase: the cumulative scores of the image go to shared memory with independent, unrolled 16-byte loads.  (Reading
        // global memory inside the first selection pass serialised ~25 dependent L2 round trips per thread: 23 us per step for
        // 50 KB per image, r2 beam timeline.)  Rows are 32-byte aligned (ldl % 8 == 0); the tail of a row is done by scalars.
        const int v4 = p.V >> 2;
        for (int r = 0; r < rows; ++r) {
            const float4* src = reinterpret_cast<const float4*>(lg + (size_t)r * p.ldl);
            float* dst = s_cache + r * p.V;
            const float tp = s_tp[r];
#pragma unroll 4
            for (int i = threadIdx.x; i < v4; i += TOPK_THREADS) {
                const float4 x = __ldg(src + i);
                dst[4 * i] = tp + x.x; dst[4 * i + 1] = tp + x.y; dst[4 * i + 2] = tp + x.z; dst[4 * i + 3] = tp + x.w;   // decoder.py:204
            }
            for (int i = 4 * v4 + threadIdx.x; i < p.V; i += TOPK_THREADS) dst[i] = tp + lg[(size_t)r * p.ldl + i];
        }
        __syncthreads();
    }
    for (int k = 0; k < n; ++k) {
        float bv = -INFINITY;
        int bi = 0x7fffffff;
        for (int i = threadIdx.x; i < total; i += TOPK_THREADS) {
            float sc;
            if (cached) {
                sc = s_cache[i];
            } else {
                const int r = i / p.V;
                sc = s_tp[r] + lg[(size_t)r * p.ldl + (i - r * p.V)];   // decoder.py:204
            }
            const bool after = (sc < last_v) || (sc == last_v && i > last_i);
            if (after && (sc > bv || (sc == bv && i < bi))) { bv = sc; bi = i; }
        }
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
        }
        if ((threadIdx.x & 31) == 0) { s_val[threadIdx.x >> 5] = bv; s_idx[threadIdx.x >> 5] = bi; }
        __syncthreads();
        if (threadIdx.x < 32) {
            bv = threadIdx.x < TOPK_THREADS / 32 ? s_val[threadIdx.x] : -INFINITY;
            bi = threadIdx.x < TOPK_THREADS / 32 ? s_idx[threadIdx.x] : 0x7fffffff;
            for (int o = 16; o > 0; o >>= 1) {
                const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
                const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
                if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
            }
            if (threadIdx.x == 0) { pick_val[k] = bv; pick_idx[k] = bi; }
        }
        __syncthreads();
        last_v = pick_val[k];
        last_i = pick_idx[k];
    }
    if (threadIdx.x == 0) {
        const size_t hb = ((size_t)p.step * p.NI + img) * p.beam;
        int nl = 0;
        int nc = p.ncomp[img];
        for (int k = 0; k < n; ++k) {
            const int flat = pick_idx[k];
            const int prev = flat / p.V, word = flat - prev * p.V;       // decoder.py:210-211
            p.cand_word[hb + k] = word;
            p.cand_parent[hb + k] = prev;
            const bool complete = (word == p.stop_a) || (word == p.stop_b);   // decoder.py:224-229 (Q15)
            if (complete) {
                p.comp_score[img * p.beam + nc] = pick_val[k];
                p.comp_step[img * p.beam + nc] = p.step;
                p.comp_cand[img * p.beam + nc] = k;
                ++nc;
            } else {
                p.slot_cand[hb + nl] = k;
                p.top_preds[img * p.beam + nl] = pick_val[k];
                p.prev_word[img * p.beam + nl] = word;
                p.parent[img * p.beam + nl] = prev;
                ++nl;
            }
        }
        p.ncomp[img] = nc;
        p.last_step[img] = p.step;
        p.last_n[img] = n;
        p.live[img] = nl;                                                 // decoder.py:238
        if (nl == 0 || p.step > p.max_steps) {                            // decoder.py:240-241, 252-253 (Q17)
            p.done[img] = 1;
            atomicSub(p.n_active, 1);
            nl = 0;
        }
        s_nl = nl;
        for (int k = 0; k < nl; ++k) s_parent[k] = p.parent[img * p.beam + k];
    }
    // h = h[prev_word_idxs[incomplete]], c = c[...] (decoder.py:246-247) into the next step's operand rows.  Source (Hn, Cn) and
    // destination (X, Cc) are different buffers, so the rows can be copied in any order.
    __syncthreads();
    const int nl = s_nl;
    const int hv = p.H * p.esize / 16, cv = p.H / 4;                      // 16-byte vectors per h row / per c row
    for (int i = threadIdx.x; i < nl * (hv + cv); i += TOPK_THREADS) {
        const int slot = i / (hv + cv), j = i - slot * (hv + cv);
        const int r = img * p.beam + slot, src = img * p.beam + s_parent[slot];
        if (j < hv)
            reinterpret_cast<uint4*>(reinterpret_cast<uint8_t*>(p.X) + ((size_t)r * p.ldx + p.hoff) * p.esize)[j] =
                reinterpret_cast<const uint4*>(reinterpret_cast<const uint8_t*>(p.Hn) + (size_t)src * p.H * p.esize)[j];
        else
            reinterpret_cast<float4*>(p.Cc + (size_t)r * p.H)[j - hv] = reinterpret_cast<const float4*>(p.Cn + (size_t)src * p.H)[j - hv];
    }
}

// every beam row of an image starts from the image's initial state
template <typename T>
__global__ void init_rows_kernel(T* __restrict__ X, int64_t ldx, int hoff, int zoff, float* __restrict__ Cc,
                                 const float* __restrict__ hc0, const T* __restrict__ abar, int write_z, int beam, int H, int D) {
    const int r = blockIdx.x, img = r / beam;
    for (int j = threadIdx.x; j < H; j += blockDim.x) {
        X[(int64_t)r * ldx + hoff + j] = from_f32<T>(hc0[(size_t)img * 2 * H + j]);
        Cc[(size_t)r * H + j] = hc0[(size_t)img * 2 * H + H + j];
    }
    if (write_z)
        for (int d = threadIdx.x; d < D; d += blockDim.x) X[(int64_t)r * ldx + zoff + d] = abar[(size_t)img * D + d];
}

struct FinalParams {
    const int* cand_word; const int* cand_parent; const int* slot_cand;
    const float* comp_score; const int* comp_step; const int* comp_cand; const int* ncomp;
    const int* last_step; const int* last_n;
    const float* alpha_hist;   // (S, R, L)
    int32_t* out_len; int32_t* out_status; int64_t* out_tokens; float* out_alphas; float* out_last_alpha; int32_t* out_live;
    int NI, beam, L, maxlen, start_token, attention;
};

__global__ void finalize_kernel(FinalParams p) {
    __shared__ int s_k[64], s_par[64];
    __shared__ int s_len;
    const int img = blockIdx.x;
    const int R = p.NI * p.beam;
    const int nc = p.ncomp[img];
    if (nc == 0) {                                                        // decoder.py:256-258 (Q18)
        const int ls = p.last_step[img], n = p.last_n[img];
        if (threadIdx.x == 0) { p.out_status[img] = 1; p.out_len[img] = 1; p.out_tokens[(size_t)img * p.maxlen] = 0; p.out_live[img] = n; }
        for (int i = threadIdx.x; i < n * p.L; i += blockDim.x) {
            const int s = i / p.L, l = i - s * p.L;
            p.out_last_alpha[((size_t)img * p.beam + s) * p.L + l] =
                p.attention ? p.alpha_hist[((size_t)ls * R + img * p.beam + s) * p.L + l] : 1.0f / (float)p.L;
        }
        return;
    }
    if (threadIdx.x == 0) {
        int best = 0;                                                     // first maximum (decoder.py:265)
        for (int i = 1; i < nc; ++i) if (p.comp_score[img * p.beam + i] > p.comp_score[img * p.beam + best]) best = i;
        const int s_end = p.comp_step[img * p.beam + best];
        int k = p.comp_cand[img * p.beam + best];
        for (int u = s_end; u >= 1; --u) {
            const size_t hb = ((size_t)u * p.NI + img) * p.beam;
            s_k[u] = k;
            const int par = p.cand_parent[hb + k];
            s_par[u] = par;
            p.out_tokens[(size_t)img * p.maxlen + u] = p.cand_word[hb + k];
            if (u > 1) k = p.slot_cand[((size_t)(u - 1) * p.NI + img) * p.beam + par];
        }
        p.out_tokens[(size_t)img * p.maxlen] = p.start_token;
        p.out_len[img] = s_end + 1;
        p.out_status[img] = 0;
        p.out_live[img] = 0;
        s_len = s_end;
    }
    __syncthreads();
    const int s_end = s_len;
    for (int l = threadIdx.x; l < p.L; l += blockDim.x) p.out_alphas[((size_t)img * p.maxlen) * p.L + l] = 1.0f;   // decoder.py:173
    for (int i = threadIdx.x; i < s_end * p.L; i += blockDim.x) {
        const int u = 1 + i / p.L, l = i % p.L;
        p.out_alphas[((size_t)img * p.maxlen + u) * p.L + l] =
            p.attention ? p.alpha_hist[((size_t)u * R + img * p.beam + s_par[u]) * p.L + l] : 1.0f / (float)p.L;
    }
}
}  // namespace beam

// comb[r, :] = p1[r, :] + p2[r, :] + emb[r, :] with emb rows strided inside X (decoder.py:155)
template <typename T>
__global__ void beam_add3_kernel(T* __restrict__ comb, const T* __restrict__ p1, const T* __restrict__ p2,
                                 const T* __restrict__ X, int64_t ldx, int H) {
    const int r = blockIdx.x;
    for (int j = threadIdx.x; j < H; j += blockDim.x)
        comb[(size_t)r * H + j] = from_f32<T>(to_f32<T>(p1[(size_t)r * H + j]) + to_f32<T>(p2[(size_t)r * H + j]) +
                                              to_f32<T>(X[(int64_t)r * ldx + j]));
}

template <typename T>
struct BeamWs {
    T *w_att, *w_uf, *w_gall, *w_init, *w_out, *w_fh, *w_fz;
    float *b_uf, *b_g, *b_init;
    T *a_c, *wa, *abar;
    float* abar_f;
    float* mean_part; // (4,NI,D) partial sums of the mean over L
    float* hc0;
    T* X;          // (R, H + D + H): [emb | z | h]
    T* Hn;         // (R, H)
    float *Cc, *Cn;
    float* UF;     // (R, H + D)
    T* Ctx;        // (R, D)
    T *P1, *P2, *Comb;
    float* logits; // (R, Vp)
    float* alpha_hist;   // (S, R, L)
    uint8_t* attn_scratch;
    bool attn_stream;
    float* top_preds;
    int64_t* prev_word;
    T *w_uf_hi, *w_uf_lo, *w_gall_hi, *w_gall_lo, *w_out_hi, *w_out_lo;      // fp32 mode: operands pre-split for the 3xTF32 GEMM
    bool presplit, merge_h;
    T* emb_t;
    float* emb_gate;
    bool emb_table;
    int *live, *done, *n_active, *parent, *cand_word, *cand_parent, *slot_cand, *comp_step, *comp_cand, *ncomp, *last_step, *last_n;
    float* comp_score;
    int64_t total;
};

extern int g_beam_emb_table, g_beam_presplit, g_beam_merge_h;
// the table costs one V x 4H x H GEMM per call and V * 4H * 4 bytes; it pays from a few decode steps on while the vocabulary is small
static inline bool beam_emb_table(const sat_beam_args& a) {
    return g_beam_emb_table && a.embedding != nullptr && (int64_t)a.V * 4 * a.H * 4 <= ((int64_t)96 << 20) && a.max_steps >= 8;
}
static inline int beam_check_args(const sat_beam_args* a) {
    SAT_REQUIRE(a != nullptr, "null args");
    SAT_REQUIRE(a->struct_bytes == (int32_t)sizeof(sat_beam_args), "sat_beam_args size mismatch: got %d, library expects %zu",
                a->struct_bytes, sizeof(sat_beam_args));
    SAT_REQUIRE(a->dtype == SAT_DTYPE_F32 || a->dtype == SAT_DTYPE_BF16, "bad dtype %d", a->dtype);
    SAT_REQUIRE(a->NI > 0 && a->L > 0 && a->D > 0 && a->H > 0 && a->V > 0, "bad dims");
    SAT_REQUIRE(a->beam >= 1 && a->beam <= beam::MAXBEAM, "beam must be in [1, %d]", beam::MAXBEAM);
    SAT_REQUIRE(a->beam <= a->V, "beam larger than the vocabulary");
    SAT_REQUIRE(a->max_steps >= 1 && a->max_steps + 1 < 64, "max_steps must be in [1, 62]");
    SAT_REQUIRE(a->D % 8 == 0 && a->H % 8 == 0, "D and H must be multiples of 8");
    return SAT_OK;
}

template <typename T>
int64_t beam_carve(const sat_beam_args& a, BeamWs<T>& w, uint8_t* base) {
    int64_t off = 0;
    auto take = [&](int64_t bytes) -> void* {
        void* p = base ? base + off : nullptr;
        off += (bytes + 255) & ~(int64_t)255;
        return p;
    };
    const int64_t NI = a.NI, L = a.L, D = a.D, H = a.H, V = a.V, R = NI * a.beam, S = a.max_steps + 2;
    // with attention on, the h share of the gate pre-activations rides on the [U; f_beta] GEMM (same operand h, 4H more output
    // columns of a launch that does not fill the GPU); the gate GEMM then contracts over [emb |] z only
    // (fp32 mode only: a 3xTF32 k-block costs ~0.7 us, so 16 fewer of them pay for the extra addend; in bf16 mode it measured 4 % slower)
    w.merge_h = a.attention && g_beam_merge_h && IsF32<T>::value;
    const int64_t ufw = H + D + (w.merge_h ? 4 * H : 0);
#define TK(field, count, type) w.field = reinterpret_cast<type*>(take((int64_t)(count) * (int64_t)sizeof(type)))
    TK(w_att, H * D, T); TK(w_uf, ufw * H, T); TK(w_gall, 4 * H * (2 * H + D), T); TK(w_init, 2 * H * D, T); TK(w_out, V * H, T);
    TK(w_fh, a.ado ? H * H : 0, T); TK(w_fz, a.ado ? H * D : 0, T);
    // embedding share of the gate pre-activations for EVERY token, once per call: emb_gate = Emb W_ih[:, :H]^T + b_ih + b_hh
    // (V x 4H fp32).  The step's gate GEMM then contracts over [z | h] only and its epilogue adds row prev_word[r].
    w.emb_table = beam_emb_table(a);
    TK(emb_t, w.emb_table ? V * H : 0, T); TK(emb_gate, w.emb_table ? V * 4 * H : 0, float);
    // fp32 mode: the weights of the three per-step GEMMs pre-split into tf32 hi / lo parts once per call (tc32::split_operand)
    w.presplit = IsF32<T>::value && g_beam_presplit && a.max_steps >= 8;
    TK(w_uf_hi, w.presplit && a.attention ? ufw * H : 0, T); TK(w_uf_lo, w.presplit && a.attention ? ufw * H : 0, T);
    TK(w_gall_hi, w.presplit ? 4 * H * (2 * H + D) : 0, T); TK(w_gall_lo, w.presplit ? 4 * H * (2 * H + D) : 0, T);
    TK(w_out_hi, w.presplit ? V * H : 0, T); TK(w_out_lo, w.presplit ? V * H : 0, T);
    TK(b_uf, ufw, float); TK(b_g, 4 * H, float); TK(b_init, 2 * H, float);
    const bool a_alias = (a.a_dtype == a.dtype) && a.a_stride_d == 1 && a.a_stride_l == D && a.a_stride_b == L * D && al16(a.img_features);
    TK(a_c, a_alias ? 0 : NI * L * D, T);
    if (a_alias) w.a_c = reinterpret_cast<T*>(const_cast<void*>(a.img_features));
    TK(wa, a.attention ? NI * L * H : 0, T); TK(abar, NI * D, T); TK(abar_f, NI * D, float); TK(mean_part, 4 * NI * D, float); TK(hc0, NI * 2 * H, float);
    TK(X, R * (2 * H + D), T); TK(Hn, R * H, T); TK(Cc, R * H, float); TK(Cn, R * H, float); TK(UF, R * ufw, float);
    TK(Ctx, R * D, T); TK(P1, a.ado ? R * H : 0, T); TK(P2, a.ado ? R * H : 0, T); TK(Comb, a.ado ? R * H : 0, T);
    TK(logits, R * ((V + 7) & ~(int64_t)7), float); TK(alpha_hist, a.attention ? S * R * L : 0, float);
    w.attn_stream = a.attention && attn_s::fwd_shape_ok(a.beam, a.L, a.D, a.H);
    TK(attn_scratch, w.attn_stream ? attn_s::scratch_bytes(a.NI, a.beam, a.D, a.H) : 0, uint8_t);
    TK(top_preds, R, float); TK(prev_word, R, int64_t);
    TK(live, NI, int); TK(done, NI, int); TK(n_active, 1, int); TK(parent, R, int);
    TK(cand_word, S * R, int); TK(cand_parent, S * R, int); TK(slot_cand, S * R, int);
    TK(comp_score, R, float); TK(comp_step, R, int); TK(comp_cand, R, int); TK(ncomp, NI, int); TK(last_step, NI, int); TK(last_n, NI, int);
#undef TK
    w.total = off;
    return off;
}

__global__ void beam_init_state_kernel(int* live, int* done, int* n_active, int* ncomp, int* last_step, int* last_n,
                                       float* top_preds, int64_t* prev_word, int NI, int beam, int64_t start) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) *n_active = NI;
    if (i < NI) { live[i] = beam; done[i] = 0; ncomp[i] = 0; last_step[i] = 0; last_n[i] = beam; }
    if (i < NI * beam) { top_preds[i] = 0.0f; prev_word[i] = start; }     // decoder.py:166-171
}

// two pinned flags + events per host thread and device for the pipelined early-exit poll
struct BeamPoll { int* flag[2]; cudaEvent_t ev[2]; bool ok; };
static inline int beam_poll_state(BeamPoll** out) {
    static thread_local BeamPoll cache[64] = {};
    int dev = 0;
    SAT_CHECK_CUDA(cudaGetDevice(&dev));
    SAT_REQUIRE(dev >= 0 && dev < 64, "device index %d out of range", dev);
    BeamPoll& b = cache[dev];
    if (!b.ok) {
        for (int i = 0; i < 2; ++i) {
            SAT_CHECK_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&b.flag[i]), sizeof(int), cudaHostAllocPortable));
            SAT_CHECK_CUDA(cudaEventCreateWithFlags(&b.ev[i], cudaEventDisableTiming));
        }
        b.ok = true;
    }
    *out = &b;
    return SAT_OK;
}

template <typename T>
int beam_search_t(const sat_beam_args& a) {
    constexpr bool PRECISE = IsF32<T>::value;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(a.stream);
    BeamWs<T> w;
    const int64_t need = beam_carve<T>(a, w, reinterpret_cast<uint8_t*>(a.workspace));
    SAT_REQUIRE(a.workspace && a.workspace_bytes >= need, "beam workspace too small: %lld < %lld", (long long)a.workspace_bytes, (long long)need);
    const int NI = a.NI, L = a.L, D = a.D, H = a.H, V = a.V, NB = a.beam, R = NI * NB;
    const int64_t xw = 2 * (int64_t)H + D;
    // ---- weights (one launch) ----
    misc::PackList<T> pl;
    if (a.attention) {
        SAT_TRY(pl.weight(w.w_att, D, a.att_W_w, D, H, D, 0, 0));
        SAT_TRY(pl.weight(w.w_uf, H, a.att_U_w, H, H, H, 0, 0));
        SAT_TRY(pl.weight(w.w_uf + (size_t)H * H, H, a.f_beta_w, H, D, H, 0, 0));
        SAT_TRY(pl.bias(w.b_uf, a.att_U_b, nullptr, H, 0));
        SAT_TRY(pl.bias(w.b_uf + H, a.f_beta_b, nullptr, D, 0));
        if (w.merge_h) {               // rows [H+D, H+D+4H) of the uf operand: W_hh, gate rows interleaved like w_gall; no bias there
            SAT_TRY(pl.weight(w.w_uf + (size_t)(H + D) * H, H, a.lstm_w_hh, H, 4 * H, H, 0, H));
            SAT_CHECK_CUDA(cudaMemsetAsync(w.b_uf + H + D, 0, (size_t)4 * H * sizeof(float), st));
        }
    }
    const int ufw = H + D + (w.merge_h ? 4 * H : 0);
    SAT_TRY(pl.weight(w.w_gall, xw, a.lstm_w_ih, H + D, 4 * H, H + D, 0, H));      // cols [0,H+D): [emb | ctx]
    SAT_TRY(pl.weight(w.w_gall + H + D, xw, a.lstm_w_hh, H, 4 * H, H, 0, H));      // cols [H+D, 2H+D): h
    SAT_TRY(pl.bias(w.b_g, a.lstm_b_ih, a.lstm_b_hh, 4 * H, H));
    SAT_TRY(pl.weight(w.w_init, D, a.init_h_w, D, H, D, 0, 0));
    SAT_TRY(pl.weight(w.w_init + (size_t)H * D, D, a.init_c_w, D, H, D, 0, 0));
    SAT_TRY(pl.bias(w.b_init, a.init_h_b, nullptr, H, 0));
    SAT_TRY(pl.bias(w.b_init + H, a.init_c_b, nullptr, H, 0));
    const float* out_b = a.ado ? a.f_out_b : a.deep_b;
    if (a.ado) {
        SAT_REQUIRE(a.f_h_w && a.f_z_w && a.f_out_w, "ado=1 needs f_h/f_z/f_out parameters");
        SAT_TRY(pl.weight(w.w_fh, H, a.f_h_w, H, H, H, 0, 0));
        SAT_TRY(pl.weight(w.w_fz, D, a.f_z_w, D, H, D, 0, 0));
        SAT_TRY(pl.weight(w.w_out, H, a.f_out_w, H, V, H, 0, 0));
    } else {
        SAT_TRY(pl.weight(w.w_out, H, a.deep_w, H, V, H, 0, 0));
    }
    if (w.emb_table) SAT_TRY(pl.weight(w.emb_t, H, a.embedding, H, V, H, 0, 0));
    SAT_TRY(pl.flush(st));
    if constexpr (IsF32<T>::value) {
        if (w.presplit) {
            if (a.attention) SAT_TRY(tc32::split_operand(st, w.w_uf_hi, w.w_uf_lo, w.w_uf, (int64_t)ufw * H));
            SAT_TRY(tc32::split_operand(st, w.w_gall_hi, w.w_gall_lo, w.w_gall, (int64_t)4 * H * xw));
            SAT_TRY(tc32::split_operand(st, w.w_out_hi, w.w_out_lo, w.w_out, (int64_t)V * H));
        }
    }
    const T* uf_hi = w.presplit ? w.w_uf_hi : nullptr;
    const T* uf_lo = w.presplit ? w.w_uf_lo : nullptr;
    const T* out_hi = w.presplit ? w.w_out_hi : nullptr;
    const T* out_lo = w.presplit ? w.w_out_lo : nullptr;
    if (w.emb_table) {
        auto e = lin<T>(V, 4 * H, w.b_g, SAT_ACT_NONE, w.emb_gate, 4 * H, nullptr, 0);
        SAT_TRY(gemm<T>(st, w.emb_t, H, true, w.w_gall, xw, true, V, 4 * H, H, e));
    }
    // ---- per-image precompute ----
    if (reinterpret_cast<const void*>(w.a_c) != a.img_features) {
        if (a.a_dtype == SAT_DTYPE_F32)
            SAT_TRY((misc::cast3d<float, T>(st, w.a_c, reinterpret_cast<const float*>(a.img_features), NI, L, D, a.a_stride_b, a.a_stride_l, a.a_stride_d)));
        else
            SAT_TRY((misc::cast3d<bf16, T>(st, w.a_c, reinterpret_cast<const bf16*>(a.img_features), NI, L, D, a.a_stride_b, a.a_stride_l, a.a_stride_d)));
    }
    SAT_TRY(misc::mean_l<T>(st, w.a_c, w.abar_f, w.abar, NI, L, D, w.mean_part, (int64_t)4 * NI * D));
    {
        auto e = lin<T>(NI, 2 * H, w.b_init, SAT_ACT_TANH, w.hc0, 2 * H, nullptr, 0);
        SAT_TRY(gemm<T>(st, w.abar, D, true, w.w_init, D, true, NI, 2 * H, D, e));
    }
    if (a.attention) {
        auto e = lin<T>(NI * L, H, a.att_W_b, SAT_ACT_NONE, nullptr, 0, w.wa, H);
        SAT_TRY(gemm<T>(st, w.a_c, D, true, w.w_att, D, true, NI * L, H, D, e));
    }
    beam::init_rows_kernel<T><<<R, 128, 0, st>>>(w.X, xw, H + D, H, w.Cc, w.hc0, w.abar, a.attention ? 0 : 1, NB, H, D);
    beam_init_state_kernel<<<sat_cdiv(R, 128), 128, 0, st>>>(w.live, w.done, w.n_active, w.ncomp, w.last_step, w.last_n, w.top_preds,
                                                             w.prev_word, NI, NB, (int64_t)a.start_token);
    SAT_CHECK_CUDA(cudaGetLastError());
    const int S = a.max_steps + 1;        // iterations 1..max_steps+1 (Q17)
    const int64_t Vp = (V + 7) & ~(int64_t)7;
    // shared-memory score cache of the top-k kernel: beam x V floats when that fits beside the static buffers
    size_t topk_smem = (size_t)NB * V * 4;
    if (topk_smem > 200 * 1024) topk_smem = 0;
    {
        static SatPerDevice configured;
        SAT_TRY(configured.once([&]() -> int {
            SAT_CHECK_CUDA(cudaFuncSetAttribute(beam::topk_step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
            return SAT_OK;
        }));
    }
    for (int step = 1; step <= S; ++step) {
        if (!w.emb_table || a.ado)      // (the advanced deep output reads the embedding itself, decoder.py:154)
            SAT_TRY(misc::gather_rows<T>(st, w.X, xw, a.embedding, H, w.prev_word, 1, R, V));              // decoder.py:183
        if (a.attention) {
            auto e = lin<T>(R, ufw, w.b_uf, SAT_ACT_NONE, w.UF, ufw, nullptr, 0);
            SAT_TRY(gemm<T>(st, w.X + H + D, xw, true, w.w_uf, H, true, R, ufw, H, e, 0, 0, 0, uf_hi, uf_lo));
            attn::FwdParams p;
            memset(&p, 0, sizeof(p));
            p.a = w.a_c; p.wa = w.wa; p.uf = w.UF; p.ld_uf = ufw; p.v = a.att_v_w; p.bv = a.att_v_b;
            p.alpha = w.alpha_hist + (size_t)step * R * L; p.ld_alpha = L;
            p.ctx = w.Ctx; p.ld_ctx = D; p.z = w.X + H; p.ld_z = xw;
            p.NI = NI; p.NB = NB; p.L = L; p.D = D; p.H = H; p.gate = a.gate;
            if (w.attn_stream) {
                p.s_part = reinterpret_cast<float*>(w.attn_scratch);
                p.static_inputs = 1;
            }
            SAT_TRY((attn_step_fwd<T, PRECISE>(p, st)));
        }
        EpiLstm<T, PRECISE> e;
        memset(&e, 0, sizeof(e));
        e.bias = w.b_g; e.c_prev = w.Cc; e.ld_cp = H; e.c_out = w.Cn; e.ld_co = H; e.h_next = w.Hn; e.ld_hn = H;
        e.M = R; e.H = H;
        const int kh = w.merge_h ? 0 : H;      // the h columns of the operand are contracted here unless the uf GEMM already did
        if (w.merge_h) { e.xg2 = w.UF + H + D; e.ld_xg2 = ufw; }
        if (w.emb_table) {              // gates = emb_gate[prev_word] + [z | h] [W_ih[:, H:] | W_hh]^T
            e.bias = nullptr; e.xg = w.emb_gate; e.ld_xg = 4 * H; e.xg_rows = w.prev_word;
            SAT_TRY(gemm<T>(st, w.X + H, xw, true, w.w_gall + H, xw, true, R, 4 * H, D + kh, e, 0, 0, 0,
                            w.presplit ? w.w_gall_hi + H : nullptr, w.presplit ? w.w_gall_lo + H : nullptr));      // decoder.py:196-197
        } else {
            SAT_TRY(gemm<T>(st, w.X, xw, true, w.w_gall, xw, true, R, 4 * H, H + D + kh, e, 0, 0, 0,
                            w.presplit ? w.w_gall_hi : nullptr, w.presplit ? w.w_gall_lo : nullptr));              // decoder.py:196-197
        }
        if (!a.ado) {
            auto eo = lin<T>(R, V, out_b, SAT_ACT_NONE, w.logits, Vp, nullptr, 0);
            SAT_TRY(gemm<T>(st, w.Hn, H, true, w.w_out, H, true, R, V, H, eo, 0, 0, 0, out_hi, out_lo));   // decoder.py:203
        } else {
            auto e1 = lin<T>(R, H, a.f_h_b, SAT_ACT_RELU, nullptr, 0, w.P1, H);
            SAT_TRY(gemm<T>(st, w.Hn, H, true, w.w_fh, H, true, R, H, H, e1));
            auto e2 = lin<T>(R, H, a.f_z_b, SAT_ACT_RELU, nullptr, 0, w.P2, H);
            if (a.attention) SAT_TRY(gemm<T>(st, w.Ctx, D, true, w.w_fz, D, true, R, H, D, e2));
            else SAT_TRY(gemm<T>(st, w.X + H, xw, true, w.w_fz, D, true, R, H, D, e2));
            // comb = P1 + P2 + emb (emb = X[:, :H], strided rows)
            beam_add3_kernel<T><<<R, 128, 0, st>>>(w.Comb, w.P1, w.P2, w.X, xw, H);
            SAT_CHECK_CUDA(cudaGetLastError());
            auto e3 = lin<T>(R, V, out_b, SAT_ACT_RELU, w.logits, Vp, nullptr, 0);
            SAT_TRY(gemm<T>(st, w.Comb, H, true, w.w_out, H, true, R, V, H, e3, 0, 0, 0, out_hi, out_lo)); // decoder.py:199-201
        }
        beam::StepParams sp;
        sp.logits = w.logits; sp.ldl = Vp; sp.top_preds = w.top_preds; sp.prev_word = w.prev_word; sp.live = w.live; sp.done = w.done;
        sp.n_active = w.n_active; sp.parent = w.parent; sp.cand_word = w.cand_word; sp.cand_parent = w.cand_parent;
        sp.slot_cand = w.slot_cand; sp.comp_score = w.comp_score; sp.comp_step = w.comp_step; sp.comp_cand = w.comp_cand;
        sp.ncomp = w.ncomp; sp.last_step = w.last_step; sp.last_n = w.last_n;
        sp.NI = NI; sp.beam = NB; sp.V = V; sp.step = step; sp.max_steps = a.max_steps;
        sp.stop_a = 1; sp.stop_b = a.bert ? 0 : 102;
        sp.cache_floats = (int)(topk_smem / 4);
        sp.X = w.X; sp.ldx = xw; sp.hoff = H + D; sp.Hn = w.Hn; sp.Cc = w.Cc; sp.Cn = w.Cn; sp.H = H; sp.esize = (int)sizeof(T);
        beam::topk_step_kernel<<<NI, beam::TOPK_THREADS, topk_smem, st>>>(sp);                    // + the state gather
        SAT_CHECK_CUDA(cudaGetLastError());
        if ((step % 8) == 0 && step < S) {     // early exit once every image has finished (no effect on results)
            // PIPELINED poll: the flag copied after step s is only looked at after the launches of steps s+1 .. s+8 have been
            // enqueued, so the host never drains the GPU (a blocking poll cost ~170 us of idle GPU each: the sync plus the time to
            // enqueue the next steps' 56 launches).  Price: up to 8 extra steps on images that are all done already -- their
            // kernels return at once (done[] / live[] guards).  Pinned flags: a pageable destination adds ~90 us of staging.
            BeamPoll* bp = nullptr;
            SAT_TRY(beam_poll_state(&bp));
            const int cur = (step / 8) & 1;
            if (step >= 16) {                  // the copy issued 8 steps ago
                SAT_CHECK_CUDA(cudaEventSynchronize(bp->ev[cur ^ 1]));
                if (*bp->flag[cur ^ 1] <= 0) break;
            }
            *bp->flag[cur] = 1;
            SAT_CHECK_CUDA(cudaMemcpyAsync(bp->flag[cur], w.n_active, sizeof(int), cudaMemcpyDeviceToHost, st));
            SAT_CHECK_CUDA(cudaEventRecord(bp->ev[cur], st));
        }
    }
    beam::FinalParams fp;
    fp.cand_word = w.cand_word; fp.cand_parent = w.cand_parent; fp.slot_cand = w.slot_cand; fp.comp_score = w.comp_score;
    fp.comp_step = w.comp_step; fp.comp_cand = w.comp_cand; fp.ncomp = w.ncomp; fp.last_step = w.last_step; fp.last_n = w.last_n;
    fp.alpha_hist = w.alpha_hist; fp.out_len = a.out_len; fp.out_status = a.out_status; fp.out_tokens = a.out_tokens;
    fp.out_alphas = a.out_alphas; fp.out_last_alpha = a.out_last_alpha; fp.out_live = a.out_live;
    fp.NI = NI; fp.beam = NB; fp.L = L; fp.maxlen = a.max_steps + 2; fp.start_token = a.start_token; fp.attention = a.attention;
    beam::finalize_kernel<<<NI, 256, 0, st>>>(fp);
    SAT_CHECK_CUDA(cudaGetLastError());
    return SAT_OK;
}
