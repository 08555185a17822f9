// GEMM epilogue functors shared by the fp32 SIMT GEMM and the bf16 tcgen05 GEMM.
// An epilogue consumes NV (multiple of 4) consecutive fp32 accumulator columns of one output row.
#pragma once
#include "sat_common.cuh"

enum { SAT_ACT_NONE = 0, SAT_ACT_TANH = 1, SAT_ACT_RELU = 2, SAT_ACT_SIGMOID = 3 };

// out = act(acc + bias[n] + addend[m,n])  -> fp32 and/or T output.  (a1,a4,a5,a8,a9 in SURVEY section 8)
template <typename T, bool PRECISE>
struct EpiLinear {
    const float* bias;       // [N] or null
    const float* addend;     // fp32 [M, ld_add] or null (may alias Cf: accumulate)
    int64_t ld_add;
    float* Cf;               // fp32 out or null
    int64_t ldcf;
    T* Ct;                   // T out or null
    int64_t ldct;
    const float* mulmask;    // optional fp32 [M, ld_mm] multiplier applied after act (e.g. relu' or dropout) or null
    int64_t ld_mm;
    int act;
    int M, N;
    int vec_ok;              // all pointers 16B aligned, all ld % 8 == 0
    int remapB, remapT;      // remapT > 0: GEMM row m = t*remapB + b is stored at output row b*remapT + t

    // L2 warm-up of what apply(m, n .. n + nc) will read (issued by the epilogue warps while the MMAs run)
    __device__ __forceinline__ void prefetch(int m, int n, int nc) const {
        if (m >= M || !addend) return;
        for (int c = 0; c < nc; c += 32) l2_prefetch(addend + (int64_t)m * ld_add + n + c);
    }

    template <int NV>
    __device__ __forceinline__ void apply(int m, int n0, const float* acc) const {
        if (m >= M) return;
        const int64_t mo = remapT > 0 ? (int64_t)(m % remapB) * remapT + m / remapB : (int64_t)m;
        uint2 pend = make_uint2(0u, 0u);
#pragma unroll
        for (int q = 0; q < NV / 4; ++q) {
            const int n = n0 + 4 * q;
            if (n >= N) return;
            float v[4] = {acc[4 * q], acc[4 * q + 1], acc[4 * q + 2], acc[4 * q + 3]};
            const bool full = vec_ok && (n + 4 <= N);
            if (full) {
                if (bias) { float4 b = *reinterpret_cast<const float4*>(bias + n); v[0] += b.x; v[1] += b.y; v[2] += b.z; v[3] += b.w; }
                if (addend) { float4 b = *reinterpret_cast<const float4*>(addend + (int64_t)m * ld_add + n); v[0] += b.x; v[1] += b.y; v[2] += b.z; v[3] += b.w; }
            } else {
#pragma unroll
                for (int i = 0; i < 4; ++i) if (n + i < N) {
                    if (bias) v[i] += bias[n + i];
                    if (addend) v[i] += addend[(int64_t)m * ld_add + n + i];
                }
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                // these activations touch only B x H sized tensors: always use the accurate versions
                if (act == SAT_ACT_TANH) v[i] = sat_tanh<true>(v[i]);
                else if (act == SAT_ACT_RELU) v[i] = fmaxf(v[i], 0.0f);
                else if (act == SAT_ACT_SIGMOID) v[i] = sat_sigmoid<true>(v[i]);
            }
            if (mulmask) {
#pragma unroll
                for (int i = 0; i < 4; ++i) if (n + i < N) v[i] *= mulmask[(int64_t)m * ld_mm + n + i];
            }
            if (full) {
                if (Cf) *reinterpret_cast<float4*>(Cf + mo * ldcf + n) = make_float4(v[0], v[1], v[2], v[3]);
                if (Ct) {
                    if (sizeof(T) == 4) {
                        *reinterpret_cast<float4*>(reinterpret_cast<float*>(Ct) + mo * ldct + n) = make_float4(v[0], v[1], v[2], v[3]);
                    } else {
                        __nv_bfloat162 lo = __floats2bfloat162_rn(v[0], v[1]), hi = __floats2bfloat162_rn(v[2], v[3]);
                        uint2 u = make_uint2(*reinterpret_cast<uint32_t*>(&lo), *reinterpret_cast<uint32_t*>(&hi));
                        bf16* dst = reinterpret_cast<bf16*>(Ct) + mo * ldct + n;
                        // 16-byte stores: an even group of four waits for its odd neighbour (n0 is a multiple of 8)
                        if (NV >= 8 && (q & 1) == 0 && n + 8 <= N) { pend = u; continue; }
                        if (NV >= 8 && (q & 1) == 1) *reinterpret_cast<uint4*>(dst - 4) = make_uint4(pend.x, pend.y, u.x, u.y);
                        else *reinterpret_cast<uint2*>(dst) = u;
                    }
                }
            } else {
#pragma unroll
                for (int i = 0; i < 4; ++i) if (n + i < N) {
                    if (Cf) Cf[mo * ldcf + n + i] = v[i];
                    if (Ct) Ct[mo * ldct + n + i] = from_f32<T>(v[i]);
                }
            }
        }
    }
};

// Epilogue of the BPTT GEMM  d[z | h_prev] = dG [W_ih[:, H:] | W_hh]  with the backward of the context gate
// (decoder.py:99-100: z = sigmoid(f_beta h) * ctx) fused in, so no separate kernel runs between this GEMM and the
// attention backward:
//   columns n <  D:  dz  ->  dctx = dz * g (+ extra), stored fp32 where dz would have gone;
//                            d(f_beta pre-activation) = dz * ctx * g (1 - g)  ->  duf[:, H + n]       (g = sigmoid(pre))
//   columns n >= D:  d h_prev, stored as is.
template <typename T, bool PRECISE>
struct EpiDzGate {
    float* out;              // fp32 [M, ld_out], D + H columns
    int64_t ld_out;
    const float* uf;         // fp32 [M, ld_uf] = [U h | f_beta h] pre-activations of this step (gate at column H + n)
    int64_t ld_uf;
    const float* ctx32;      // fp32 [M, ld_c32] un-gated context saved by the forward attention kernel
    int64_t ld_c32;
    const float* extra;      // optional fp32 [M, ld_ex]: gradient reaching the UN-gated context (ado, decoder.py:120)
    int64_t ld_ex;
    T* duf;                  // [M, ld_duf]: d[U h | f_beta h]; this epilogue writes columns H .. H + D
    int64_t ld_duf;
    int M, D, H, gate;

    __device__ __forceinline__ void prefetch(int m, int n, int nc) const {
        if (m >= M || n >= D) return;
        for (int c = 0; c < nc; c += 32) {
            if (gate) { l2_prefetch(uf + (int64_t)m * ld_uf + H + n + c); l2_prefetch(ctx32 + (int64_t)m * ld_c32 + n + c); }
            if (extra) l2_prefetch(extra + (int64_t)m * ld_ex + n + c);
        }
    }

    template <int NV>
    __device__ __forceinline__ void apply(int m, int n0, const float* acc) const {
        if (m >= M) return;
#pragma unroll
        for (int q = 0; q < NV / 4; ++q) {
            const int n = n0 + 4 * q;
            if (n >= D + H) return;
            float v[4] = {acc[4 * q], acc[4 * q + 1], acc[4 * q + 2], acc[4 * q + 3]};
            if (n < D) {
                float dfb[4] = {0.0f, 0.0f, 0.0f, 0.0f};
                if (gate) {
                    const float4 gp = *reinterpret_cast<const float4*>(uf + (int64_t)m * ld_uf + H + n);
                    const float4 cx = *reinterpret_cast<const float4*>(ctx32 + (int64_t)m * ld_c32 + n);
                    const float g[4] = {sat_sigmoid<PRECISE>(gp.x), sat_sigmoid<PRECISE>(gp.y), sat_sigmoid<PRECISE>(gp.z), sat_sigmoid<PRECISE>(gp.w)};
                    const float c[4] = {cx.x, cx.y, cx.z, cx.w};
#pragma unroll
                    for (int i = 0; i < 4; ++i) { dfb[i] = v[i] * c[i] * g[i] * (1.0f - g[i]); v[i] *= g[i]; }
                }
                store4<T>(duf + (int64_t)m * ld_duf + H + n, dfb);
                if (extra) {
                    const float4 e = *reinterpret_cast<const float4*>(extra + (int64_t)m * ld_ex + n);
                    v[0] += e.x; v[1] += e.y; v[2] += e.z; v[3] += e.w;
                }
            }
            *reinterpret_cast<float4*>(out + (int64_t)m * ld_out + n) = make_float4(v[0], v[1], v[2], v[3]);
        }
    }
};

// Epilogue of the BPTT GEMM  d h_{t-1} = dZH[:, D:] + [dUh | dfb] [U ; f_beta]  with the LSTM cell backward of step t-1 fused in
// (the pointwise kernel misc::lstm_bwd_kernel, same arithmetic): column n is hidden unit n, so the epilogue turns d h_{t-1} of
// a (row, unit) straight into the four gate gradients of step t-1 and the carried d c, and one launch per BPTT step goes away.
//   dh   = acc + addend + dhd * mask            (dhd = gradient from the output layer at step t-1, through dropout)
//   d_o  = dh tanh(c) o (1-o) ;  dc = dh o (1 - tanh(c)^2) + dc_carry
//   d_i  = dc g i (1-i) ;  d_f = dc c_prev f (1-f) ;  d_g = dc i (1-g^2) ;  dc_carry <- dc f
template <typename T, bool PRECISE>
struct EpiDhCell {
    const float* addend;     // fp32 [M, ld_add]: direct d h_prev of this step (dZH[:, D:])
    int64_t ld_add;
    const T* gates;          // [M, ld_g] activated gates of step t-1 (interleaved i,f,g,o per unit)
    int64_t ld_g;
    const float* c_prev;     // fp32 [M, ld_cp] cell state entering step t-1
    int64_t ld_cp;
    const float* c_new;      // fp32 [M, ld_cn] cell state leaving step t-1
    int64_t ld_cn;
    const float* dhd;        // fp32 [M, ld_dhd] or null
    int64_t ld_dhd;
    const float* mask;       // fp32 [M, ld_mask] scaled keep mask or null
    int64_t ld_mask;
    float* dc_carry;         // fp32 [M, ld_dc], read-modify-write
    int64_t ld_dc;
    T* dG;                   // [M, ld_dg] gate gradients of step t-1 (interleaved)
    int64_t ld_dg;
    int M, H;

    __device__ __forceinline__ void prefetch(int m, int n, int nc) const {
        if (m >= M) return;
        for (int c = 0; c < 4 * nc; c += 128 / (int)sizeof(T)) l2_prefetch(gates + (int64_t)m * ld_g + 4 * n + c);
        l2_prefetch(c_prev + (int64_t)m * ld_cp + n);
        l2_prefetch(c_new + (int64_t)m * ld_cn + n);
        if (dhd) l2_prefetch(dhd + (int64_t)m * ld_dhd + n);
        if (mask) l2_prefetch(mask + (int64_t)m * ld_mask + n);
    }

    template <int NV>
    __device__ __forceinline__ void apply(int m, int n0, const float* acc) const {
        if (m >= M) return;
#pragma unroll
        for (int q = 0; q < NV / 4; ++q) {
            const int n = n0 + 4 * q;
            if (n >= H) return;                       // H is a multiple of 8: groups of four units are all-in or all-out
            float dh[4], cp[4], cn[4], dcc[4];
            load4<float>(addend + (int64_t)m * ld_add + n, dh);
#pragma unroll
            for (int i = 0; i < 4; ++i) dh[i] += acc[4 * q + i];
            if (dhd) {
                float d[4], mk[4] = {1.0f, 1.0f, 1.0f, 1.0f};
                load4<float>(dhd + (int64_t)m * ld_dhd + n, d);
                if (mask) load4<float>(mask + (int64_t)m * ld_mask + n, mk);
#pragma unroll
                for (int i = 0; i < 4; ++i) dh[i] += d[i] * mk[i];
            }
            load4<float>(c_prev + (int64_t)m * ld_cp + n, cp);
            load4<float>(c_new + (int64_t)m * ld_cn + n, cn);
            load4<float>(dc_carry + (int64_t)m * ld_dc + n, dcc);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                float g[4];
                load4<T>(gates + (int64_t)m * ld_g + 4 * (n + i), g);
                const float gi = g[0], gf = g[1], gg = g[2], go = g[3];
                const float tc = tanhf(cn[i]);
                const float d_o = dh[i] * tc * go * (1.0f - go);
                const float dc = dh[i] * go * (1.0f - tc * tc) + dcc[i];
                const float o4[4] = {dc * gg * gi * (1.0f - gi), dc * cp[i] * gf * (1.0f - gf), dc * gi * (1.0f - gg * gg), d_o};
                dcc[i] = dc * gf;
                store4<T>(dG + (int64_t)m * ld_dg + 4 * (n + i), o4);
            }
            *reinterpret_cast<float4*>(dc_carry + (int64_t)m * ld_dc + n) = make_float4(dcc[0], dcc[1], dcc[2], dcc[3]);
        }
    }
};

// LSTM cell epilogue (a7: decoder.py:115 nn.LSTMCell).  The gate GEMM's N axis is gate-interleaved:
// column n = 4*j + g with g in (i,f,g,o) for hidden unit j, so one group of 4 columns is one unit.
template <typename T, bool PRECISE>
struct EpiLstm {
    const float* xg;         // fp32 [M, ld_xg] pre-activation addend (interleaved) or null
    int64_t ld_xg;
    const int64_t* xg_rows;  // null: row m of xg belongs to GEMM row m; else row xg_rows[m] (beam search: a table indexed by token id)
    const float* xg2;        // fp32 [M, ld_xg2] second pre-activation addend (interleaved) or null (beam search: the h share of the gates)
    int64_t ld_xg2;
    const float* bias;       // fp32 [4H] interleaved or null
    const float* c_prev;     // fp32 [M, ld_cp]
    int64_t ld_cp;
    float* c_out;            // fp32 [M, ld_co]
    int64_t ld_co;
    T* h_next;               // [M, ld_hn] un-dropped h for the next step's GEMM operand, or null
    int64_t ld_hn;
    T* h_out;                // [M, ld_ho] h * mask (input of the deep-output layer), or null
    int64_t ld_ho;
    const float* mask;       // fp32 [M, ld_mask] scaled keep mask or null (eval)
    int64_t ld_mask;
    T* gates;                // [M, ld_g] activated gates (interleaved) saved for backward, or null
    int64_t ld_g;
    int M, H;

    // one hidden unit: 4 gate pre-activations -> (c, h, activated gates)
    __device__ __forceinline__ void cell(float* p, float cp, float& c, float& h) const {
        p[0] = sat_sigmoid<true>(p[0]);
        p[1] = sat_sigmoid<true>(p[1]);
        p[2] = sat_tanh<true>(p[2]);
        p[3] = sat_sigmoid<true>(p[3]);
        c = p[1] * cp + p[0] * p[2];
        h = p[3] * sat_tanh<true>(c);
    }

    // the epilogue's inputs (this step's Xg rows, the previous cell state, the dropout mask) were last touched before the
    // attention kernel streamed 257 MB through L2: request them while the MMAs run instead of after the accumulator is ready
    __device__ __forceinline__ void prefetch(int m, int n, int nc) const {
        if (m >= M) return;
        if (xg && !xg_rows) for (int c = 0; c < nc; c += 32) l2_prefetch(xg + (int64_t)m * ld_xg + n + c);
        l2_prefetch(c_prev + (int64_t)m * ld_cp + (n >> 2));
        if (mask) l2_prefetch(mask + (int64_t)m * ld_mask + (n >> 2));
    }

    template <int NV>
    __device__ __forceinline__ void apply(int m, int n0, const float* acc) const {
        if (m >= M) return;
        const int64_t xrow = (xg && xg_rows) ? xg_rows[m] : (int64_t)m;
        if constexpr (NV >= 16) {
            // the tensor-core GEMM hands one thread NV consecutive columns of ONE row = NV/4 consecutive units: work on
            // groups of 4 units so that every global access is a 16-byte (fp32) or 8-byte (bf16) vector
#pragma unroll
            for (int u0 = 0; u0 < NV / 4; u0 += 4) {
                const int j0 = (n0 >> 2) + u0;
                if (j0 >= H) return;
                const float4 cp4 = *reinterpret_cast<const float4*>(c_prev + (int64_t)m * ld_cp + j0);
                const float cp[4] = {cp4.x, cp4.y, cp4.z, cp4.w};
                float c[4], h[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int n = n0 + 4 * (u0 + u);
                    float p[4] = {acc[4 * (u0 + u)], acc[4 * (u0 + u) + 1], acc[4 * (u0 + u) + 2], acc[4 * (u0 + u) + 3]};
                    if (xg) { float4 x = *reinterpret_cast<const float4*>(xg + xrow * ld_xg + n); p[0] += x.x; p[1] += x.y; p[2] += x.z; p[3] += x.w; }
                    if (xg2) { float4 x = *reinterpret_cast<const float4*>(xg2 + (int64_t)m * ld_xg2 + n); p[0] += x.x; p[1] += x.y; p[2] += x.z; p[3] += x.w; }
                    if (bias) { float4 x = *reinterpret_cast<const float4*>(bias + n); p[0] += x.x; p[1] += x.y; p[2] += x.z; p[3] += x.w; }
                    cell(p, cp[u], c[u], h[u]);
                    if (gates) store4<T>(gates + (int64_t)m * ld_g + n, p);
                }
                *reinterpret_cast<float4*>(c_out + (int64_t)m * ld_co + j0) = make_float4(c[0], c[1], c[2], c[3]);
                if (h_next) store4<T>(h_next + (int64_t)m * ld_hn + j0, h);
                if (h_out) {
                    if (mask) {
                        const float4 mk = *reinterpret_cast<const float4*>(mask + (int64_t)m * ld_mask + j0);
                        h[0] *= mk.x; h[1] *= mk.y; h[2] *= mk.z; h[3] *= mk.w;
                    }
                    store4<T>(h_out + (int64_t)m * ld_ho + j0, h);
                }
            }
        } else {
#pragma unroll
            for (int q = 0; q < NV / 4; ++q) {
                const int n = n0 + 4 * q;
                const int j = n >> 2;
                if (j >= H) return;
                float p[4] = {acc[4 * q], acc[4 * q + 1], acc[4 * q + 2], acc[4 * q + 3]};
                if (xg) { float4 x = *reinterpret_cast<const float4*>(xg + xrow * ld_xg + n); p[0] += x.x; p[1] += x.y; p[2] += x.z; p[3] += x.w; }
                if (xg2) { float4 x = *reinterpret_cast<const float4*>(xg2 + (int64_t)m * ld_xg2 + n); p[0] += x.x; p[1] += x.y; p[2] += x.z; p[3] += x.w; }
                if (bias) { float4 x = *reinterpret_cast<const float4*>(bias + n); p[0] += x.x; p[1] += x.y; p[2] += x.z; p[3] += x.w; }
                float c, h;
                cell(p, c_prev[(int64_t)m * ld_cp + j], c, h);
                c_out[(int64_t)m * ld_co + j] = c;
                if (h_next) h_next[(int64_t)m * ld_hn + j] = from_f32<T>(h);
                if (h_out) {
                    const float mk = mask ? mask[(int64_t)m * ld_mask + j] : 1.0f;
                    h_out[(int64_t)m * ld_ho + j] = from_f32<T>(h * mk);
                }
                if (gates) store4<T>(gates + (int64_t)m * ld_g + n, p);
            }
        }
    }
};
