// Fused soft-attention decode step (attention.py:14-21 + gate decoder.py:99-100), forward and backward.
//
// One thread-block CLUSTER per image: CTA `rank` owns the L-chunk [rank*Lc, rank*Lc+Lc) of the image's L
// locations.  Every byte the CTA needs from HBM -- its contiguous chunk of encoder features a[img, l0:l1, :]
// and of the hoisted projection Wa[img, l0:l1, :] -- is requested at kernel entry by the TMA engine
// (cp.async.bulk, mbarrier completion), so each resident CTA keeps its whole working set in flight while the
// other CTAs of the SM compute.  Forward, per CTA:
//   (1) scores e_l = v . tanh(Wa[l,:] + Uh) + b_v  (warp per location, warp-shuffle reduction),
//   (2) a chunk-local (online) softmax gives p_l = exp(e_l - m_rank) and s_rank,
//   (3) the un-normalised partial context sum_l p_l a_l is accumulated from shared memory,
//   (4) (m, s, partial ctx) are exchanged through distributed shared memory; every CTA finalises the alphas of
//       its chunk and a D-slice of the context, applies the sigmoid(f_beta h) gate and writes alpha (fp32), ctx
//       and the gated context.
// Each element of `a` and `Wa` is read from HBM exactly once per step: L*(H+D)*sizeof(T) bytes per image.
// NB > 1 rows per image (beam search) share the staged features.
#pragma once
#include "sat_common.cuh"

namespace attn {
constexpr int THREADS = 256;
constexpr int NWARPS = THREADS / 32;
constexpr int MAXC = 16;       // max cluster size (16 needs the non-portable opt-in)
constexpr int GMAX = 8;        // max L-groups in the context phase
constexpr uint32_t BULK_PIECE = 16384;

struct FwdParams {
    const void* a;        // (NI, L, D) T
    const void* wa;       // (NI, L, H) T
    const float* uf;      // (NR, ld_uf) fp32: [0,H) = U h + b_U ; [H,H+D) = f_beta h + b (pre-sigmoid)
    int64_t ld_uf;
    const float* v;       // (H) attention.v.weight
    const float* bv;      // (1) attention.v.bias
    float* alpha;         // row r -> alpha + r*ld_alpha (L floats)
    int64_t ld_alpha;
    void* ctx;            // T (NR, ld_ctx) un-gated context, or null
    int64_t ld_ctx;
    void* z;              // T (NR, ld_z) gated context, or null
    int64_t ld_z;
    int NI, NB, L, D, H;
    int gate;             // apply sigmoid gate
    int C, Lc, Ds, G;
    // streaming kernels only (sat_attn_stream.cuh)
    float* ctx32;         // fp32 (NR, ld_c32) un-gated context saved for the backward pass, or null
    int64_t ld_c32;
    float* s_part;        // cross-CTA partial records (scratch)
    int static_inputs;    // a / wa were NOT written by the kernel launched just before this one: the producers may
                          // start streaming them before the programmatic-dependent-launch wait
    int max_ctas;         // > 0: the grid uses at most this many CTAs (leaves SMs to kernels running beside it)
};

struct SmemLayout {
    size_t a_off, wa_off, uh_off, v_off, e_off, ms_off, ctx_off, f_off, bar_off, total;
    int lcpad;
};

// one elected thread: stream `bytes` from global to shared in pieces, all completing on `bar`
__device__ __forceinline__ void bulk_load_region(void* sdst, const void* gsrc, size_t bytes, uint64_t* bar) {
    if (bytes == 0) { mbar_arrive(bar); return; }
    mbar_expect_tx(bar, (uint32_t)bytes);
    for (size_t o = 0; o < bytes; o += BULK_PIECE) {
        const uint32_t n = (uint32_t)((bytes - o) < BULK_PIECE ? (bytes - o) : BULK_PIECE);
        bulk_g2s(reinterpret_cast<uint8_t*>(sdst) + o, reinterpret_cast<const uint8_t*>(gsrc) + o, n, bar);
    }
}

template <typename T>
inline SmemLayout fwd_layout(int NBG, int NB, int Lc, int D, int H, int G) {
    SmemLayout s;
    size_t o = 0;
    const size_t a_bytes = ((size_t)Lc * D * sizeof(T) + 127) & ~(size_t)127;
    const size_t ctx_bytes = (size_t)G * NBG * D * 4;
    // single beam group: the partial-context buffer may reuse the staged-feature region
    const bool alias = (NB <= NBG) && (ctx_bytes <= a_bytes);
    s.a_off = o; o += a_bytes;
    s.wa_off = o; o += ((size_t)Lc * H * sizeof(T) + 127) & ~(size_t)127;
    s.uh_off = o; o += (size_t)NBG * H * 4;
    s.v_off = o; o += (size_t)H * 4;
    s.lcpad = (Lc + 3) & ~3;
    s.e_off = o; o += (size_t)NBG * s.lcpad * 4;
    s.ms_off = o; o += 64;
    if (alias) s.ctx_off = s.a_off; else { s.ctx_off = o; o += ctx_bytes; }
    s.f_off = o; o += (size_t)MAXC * 8 * 4;
    o = (o + 15) & ~(size_t)15;
    s.bar_off = o; o += 2 * 8;
    s.total = o;
    return s;
}

template <typename T, int NBG, bool PRECISE>
__global__ void __launch_bounds__(THREADS)
fwd_kernel(FwdParams p, SmemLayout sl) {
    constexpr int VEC = Vec16<T>::N;
    extern __shared__ __align__(128) uint8_t smem[];
    T* a_s = reinterpret_cast<T*>(smem + sl.a_off);
    T* wa_s = reinterpret_cast<T*>(smem + sl.wa_off);
    float* uh_s = reinterpret_cast<float*>(smem + sl.uh_off);
    float* v_s = reinterpret_cast<float*>(smem + sl.v_off);
    float* e_s = reinterpret_cast<float*>(smem + sl.e_off);
    float* ms_s = reinterpret_cast<float*>(smem + sl.ms_off);     // [nb][2]
    float* ctx_s = reinterpret_cast<float*>(smem + sl.ctx_off);   // [g][nb][D]
    float* f_s = reinterpret_cast<float*>(smem + sl.f_off);       // [rank][8]
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + sl.bar_off);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int rank = blockIdx.x;            // == cluster CTA rank (gridDim.x == cluster size)
    const int img = blockIdx.y;
    const int L = p.L, D = p.D, H = p.H, C = p.C;
    const int l0 = min(L, rank * p.Lc), l1 = min(L, l0 + p.Lc);
    const int Lr = l1 - l0;
    const int lcpad = sl.lcpad;

    // (0) put this CTA's whole HBM working set in flight
    if (tid == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        mbar_fence_init();
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        bulk_load_region(wa_s, reinterpret_cast<const T*>(p.wa) + ((size_t)img * L + l0) * H, (size_t)Lr * H * sizeof(T), &bars[0]);
        bulk_load_region(a_s, reinterpret_cast<const T*>(p.a) + ((size_t)img * L + l0) * D, (size_t)Lr * D * sizeof(T), &bars[1]);
    }
    for (int i = tid; i < H; i += THREADS) v_s[i] = p.v[i];
    const float bv = p.bv[0];
    const int HV = H / VEC;
    const int nchunks = D / VEC;
    bool staged = false;

    for (int bg = 0; bg * NBG < p.NB; ++bg) {
        const int row_base = img * p.NB + bg * NBG;
        const int nb_valid = min(NBG, p.NB - bg * NBG);
        for (int i = tid; i < NBG * H; i += THREADS) {
            const int nb = i / H, h = i - nb * H;
            uh_s[i] = (nb < nb_valid) ? p.uf[(int64_t)(row_base + nb) * p.ld_uf + h] : 0.0f;
        }
        __syncthreads();
        if (!staged) mbar_wait(&bars[0], 0);

        // (1) scores: warp per location, lanes stride over 16-byte chunks of the staged Wa row
        for (int li = warp; li < Lr; li += NWARPS) {
            const T* wrow = wa_s + (size_t)li * H;
            float s[NBG];
#pragma unroll
            for (int nb = 0; nb < NBG; ++nb) s[nb] = 0.0f;
            for (int c = lane; c < HV; c += 32) {
                const uint4 u = *reinterpret_cast<const uint4*>(wrow + (size_t)c * VEC);
                float w[VEC];
                Vec16<T>::unpack(u, w);
                const float* vv = v_s + c * VEC;
#pragma unroll
                for (int nb = 0; nb < NBG; ++nb) {
                    const float* uu = uh_s + nb * H + c * VEC;
#pragma unroll
                    for (int i = 0; i < VEC; ++i) s[nb] = fmaf(vv[i], attn_tanh<PRECISE>(w[i] + uu[i]), s[nb]);
                }
            }
#pragma unroll
            for (int nb = 0; nb < NBG; ++nb) {
                const float t = warp_sum(s[nb]);
                if (lane == 0) e_s[nb * lcpad + li] = t + bv;
            }
        }
        __syncthreads();

        // (2) chunk-local softmax statistics: warp nb handles beam row nb
        for (int nb = warp; nb < NBG; nb += NWARPS) {
            float m = -INFINITY;
            for (int li = lane; li < Lr; li += 32) m = fmaxf(m, e_s[nb * lcpad + li]);
            m = warp_max(m);
            float sum = 0.0f;
            for (int li = lane; li < Lr; li += 32) {
                const float pe = sat_exp<PRECISE>(e_s[nb * lcpad + li] - m);
                e_s[nb * lcpad + li] = pe;
                sum += pe;
            }
            sum = warp_sum(sum);
            if (lane == 0) { ms_s[nb * 2] = m; ms_s[nb * 2 + 1] = sum; }
        }
        __syncthreads();

        // (3) partial context from the staged features
        if (!staged) { mbar_wait(&bars[1], 0); staged = true; }
        {
            // thread -> (L-group g, 16-byte chunk ch); for wide D one thread owns up to KDMAX chunks
            const int G = p.G;
            constexpr int KDMAX = 2;
            float acc[KDMAX][NBG][VEC];
#pragma unroll
            for (int kd = 0; kd < KDMAX; ++kd)
#pragma unroll
                for (int nb = 0; nb < NBG; ++nb)
#pragma unroll
                    for (int i = 0; i < VEC; ++i) acc[kd][nb][i] = 0.0f;
            const bool wide = nchunks >= THREADS;
            const int g = wide ? 0 : tid / nchunks;
            const int ch0 = wide ? tid : tid - g * nchunks;
            const int lstep = wide ? 1 : G;
            if (g < G) {
#pragma unroll
                for (int kd = 0; kd < KDMAX; ++kd) {
                    const int ch = ch0 + kd * THREADS;
                    if (ch < nchunks && (wide || kd == 0)) {
#pragma unroll 4
                        for (int li = g; li < Lr; li += lstep) {
                            const uint4 u = *reinterpret_cast<const uint4*>(a_s + (size_t)li * D + ch * VEC);
                            float f[VEC];
                            Vec16<T>::unpack(u, f);
#pragma unroll
                            for (int nb = 0; nb < NBG; ++nb) {
                                const float pe = e_s[nb * lcpad + li];
#pragma unroll
                                for (int i = 0; i < VEC; ++i) acc[kd][nb][i] = fmaf(pe, f[i], acc[kd][nb][i]);
                            }
                        }
                    }
                }
            }
            __syncthreads();        // ctx_s may alias the staged features: all reads of a_s are done
            if (g < G) {
#pragma unroll
                for (int kd = 0; kd < KDMAX; ++kd) {
                    const int ch = ch0 + kd * THREADS;
                    if (ch < nchunks && (wide || kd == 0)) {
#pragma unroll
                        for (int nb = 0; nb < NBG; ++nb)
#pragma unroll
                            for (int i4 = 0; i4 < VEC / 4; ++i4)
                                *reinterpret_cast<float4*>(&ctx_s[((size_t)g * NBG + nb) * D + ch * VEC + 4 * i4]) =
                                    make_float4(acc[kd][nb][4 * i4], acc[kd][nb][4 * i4 + 1], acc[kd][nb][4 * i4 + 2], acc[kd][nb][4 * i4 + 3]);
                    }
                }
            }
            if (G > 1) {
                __syncthreads();
                for (int i = tid; i < NBG * D; i += THREADS) {
                    float t = ctx_s[i];
                    for (int gg = 1; gg < G; ++gg) t += ctx_s[(size_t)gg * NBG * D + i];
                    ctx_s[i] = t;
                }
            }
        }
        // (4) cluster exchange
        cluster_sync_all();
        if (tid < NBG) {
            float mr[MAXC], sr[MAXC];
            float M = -INFINITY;
#pragma unroll
            for (int r = 0; r < MAXC; ++r) if (r < C) {
                const uint32_t ad = dsmem_addr(ms_s + tid * 2, r);
                mr[r] = dsmem_ld_f32(ad);
                sr[r] = dsmem_ld_f32(ad + 4);
                M = fmaxf(M, mr[r]);
            }
            float S = 0.0f;
#pragma unroll
            for (int r = 0; r < MAXC; ++r) if (r < C) { mr[r] = sat_exp<PRECISE>(mr[r] - M); S = fmaf(sr[r], mr[r], S); }
            const float inv = 1.0f / S;
#pragma unroll
            for (int r = 0; r < MAXC; ++r) if (r < C) f_s[r * 8 + tid] = mr[r] * inv;
        }
        __syncthreads();
        for (int i = tid; i < nb_valid * Lr; i += THREADS) {
            const int nb = i / Lr, li = i - nb * Lr;
            p.alpha[(int64_t)(row_base + nb) * p.ld_alpha + l0 + li] = e_s[nb * lcpad + li] * f_s[rank * 8 + nb];
        }
        {
            // this CTA finalises the D-slice [d0, d1) (Ds is a multiple of 4): 4 columns per thread, float4 DSMEM reads
            const int d0 = min(D, rank * p.Ds), d1 = min(D, d0 + p.Ds);
            const int nq = (d1 - d0) / 4;
            for (int i = tid; i < nb_valid * nq; i += THREADS) {
                const int nb = i / nq, d = d0 + 4 * (i - nb * nq);
                float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
                for (int r = 0; r < C; ++r) {
                    const float f = f_s[r * 8 + nb];
                    const float4 x = dsmem_ld_f32x4(dsmem_addr(ctx_s + nb * D + d, r));
                    t.x = fmaf(f, x.x, t.x); t.y = fmaf(f, x.y, t.y); t.z = fmaf(f, x.z, t.z); t.w = fmaf(f, x.w, t.w);
                }
                const int64_t row = row_base + nb;
                const float tv[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    if (p.ctx) reinterpret_cast<T*>(p.ctx)[row * p.ld_ctx + d + k] = from_f32<T>(tv[k]);
                    if (p.z) {
                        const float g = p.gate ? sat_sigmoid<PRECISE>(p.uf[row * p.ld_uf + H + d + k]) : 1.0f;
                        reinterpret_cast<T*>(p.z)[row * p.ld_z + d + k] = from_f32<T>(tv[k] * g);
                    }
                }
            }
        }
        cluster_sync_all();   // peers finished reading this CTA's smem; also orders reuse in the next beam group
    }
}

// ------------------------------------------------------------------------------------------------
// backward of one attention step (NB == 1).  Per CTA (same cluster/chunk decomposition as forward):
//   (0) a-chunk and Wa-chunk staged by TMA; the old values of this CTA's slice of the fp32 accumulator
//       sum_t dpre (B,L,H) are prefetched into registers; dctx = dz*gate (+extra) -> shared
//   (1) dalpha_l = dctx . a_l (+ external grad)      (warp per location, from shared memory)
//   (2) cluster exchange of sum_l alpha_l dalpha_l -> de_l = alpha_l (dalpha_l - S)     (softmax backward)
//   (3) pass over staged Wa: dpre = de_l v (1 - tanh^2); accumulator += dpre; dUh += dpre; dv += de_l tanh
//   (4) cluster reduction of dUh / dv partials through DSMEM; each CTA writes an H-slice.
// ------------------------------------------------------------------------------------------------
struct BwdParams {
    const void* a;          // (B, L, D) T
    const void* wa;         // (B, L, H) T
    const float* uf;        // saved (B, ld_uf) [Uh | f_beta pre]
    int64_t ld_uf;
    const float* v;         // (H)
    const float* alpha;     // saved alpha rows, row b -> alpha + b*ld_alpha
    int64_t ld_alpha;
    const void* ctx;        // saved un-gated ctx T (B, ld_ctx)
    int64_t ld_ctx;
    const float* dz;        // grad wrt gated context (B, ld_dz) fp32
    int64_t ld_dz;
    const float* dctx_extra;  // grad wrt un-gated context (ado path) fp32 or null
    int64_t ld_dce;
    const float* galpha;    // external grad wrt alpha rows or null
    int64_t ld_ga;
    void* duf;              // T (B, ld_duf): [dUh (H) | d f_beta pre (D)]
    int64_t ld_duf;
    float* dwa_acc;         // (B, L, H) fp32 accumulated over time
    float* dv_acc;          // (B, H) fp32
    float* dbv_acc;         // (B)   fp32
    float* da_acc;          // (B, L, D) fp32 or null
    void* dpre_out;         // T (B, L, H): this step's dpre written once (streaming kernel); null -> accumulate into dwa_acc
    float* de_out;          // streaming kernel: (B, L) fp32 score gradients de of this step (or null).  With de_out set dpre_out may
                            // be null: the caller rebuilds sum_t dpre from de / uf / wa afterwards (misc::dpre_sum)
    const float* ctx32;     // streaming kernel: fp32 (B, ld_c32) un-gated context saved by the forward kernel
    int64_t ld_c32;
    float* s_part;          // streaming kernel scratch (see FwdParams)
    float* s_dctx;          // streaming kernel scratch: (B, D) fp32 gated upstream gradient (when !dz_is_dctx)
    int dz_is_dctx;         // streaming kernel: dz already holds dctx = dz*gate + extra and d f_beta is already written
    int static_inputs;      // see FwdParams
    int first;              // first backward step: overwrite the accumulators
    int B, L, D, H, gate;
    int C, Lc, Ds, Hs, G2;
    int max_ctas;         // > 0: the grid uses at most this many CTAs
};

struct BwdSmem {
    size_t a_off, wa_off, dctx_off, uh_off, v_off, dal_off, red_off, misc_off, bar_off, total;
};

template <typename T>
inline BwdSmem bwd_layout(int Lc, int D, int H, int G2) {
    BwdSmem s;
    size_t o = 0;
    const size_t a_bytes = ((size_t)Lc * D * sizeof(T) + 127) & ~(size_t)127;
    s.a_off = o; o += a_bytes;
    s.wa_off = o; o += ((size_t)Lc * H * sizeof(T) + 127) & ~(size_t)127;
    s.dctx_off = o; o += ((size_t)D * sizeof(T) + 15) & ~(size_t)15;
    s.uh_off = o; o += (size_t)H * 4;
    s.v_off = o; o += (size_t)H * 4;
    s.dal_off = o; o += (size_t)((Lc + 3) & ~3) * 4 * 2;   // dalpha/de and alpha
    // the [G2][2][H] dUh/dv partial buffer is only live after the features are consumed: reuse their region
    const size_t red_bytes = (size_t)G2 * 2 * H * 4;
    if (red_bytes <= a_bytes) s.red_off = s.a_off; else { s.red_off = o; o += red_bytes; }
    s.misc_off = o; o += 64;
    o = (o + 15) & ~(size_t)15;
    s.bar_off = o; o += 2 * 8;
    s.total = o;
    return s;
}

template <typename T, bool PRECISE>
__global__ void __launch_bounds__(THREADS, 3)
bwd_kernel(BwdParams p, BwdSmem sl) {
    constexpr int VEC = Vec16<T>::N;
    constexpr int RP = 4;                   // accumulator rows prefetched per thread
    extern __shared__ __align__(128) uint8_t smem[];
    T* a_s = reinterpret_cast<T*>(smem + sl.a_off);
    T* wa_s = reinterpret_cast<T*>(smem + sl.wa_off);
    T* dctx_s = reinterpret_cast<T*>(smem + sl.dctx_off);
    float* uh_s = reinterpret_cast<float*>(smem + sl.uh_off);
    float* v_s = reinterpret_cast<float*>(smem + sl.v_off);
    float* de_s = reinterpret_cast<float*>(smem + sl.dal_off);
    const int lcpad = (p.Lc + 3) & ~3;
    float* al_s = de_s + lcpad;
    float* red_s = reinterpret_cast<float*>(smem + sl.red_off);   // [g][2][H]: dUh partial, dv partial
    float* misc_s = reinterpret_cast<float*>(smem + sl.misc_off); // [0]=sum alpha*dalpha, [1]=sum de, [2]=S, [4..12) scratch
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + sl.bar_off);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int rank = blockIdx.x, b = blockIdx.y;
    const int L = p.L, D = p.D, H = p.H, C = p.C;
    const int l0 = min(L, rank * p.Lc), l1 = min(L, l0 + p.Lc);
    const int Lr = l1 - l0;

    if (tid == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        mbar_fence_init();
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        bulk_load_region(a_s, reinterpret_cast<const T*>(p.a) + ((size_t)b * L + l0) * D, (size_t)Lr * D * sizeof(T), &bars[0]);
        bulk_load_region(wa_s, reinterpret_cast<const T*>(p.wa) + ((size_t)b * L + l0) * H, (size_t)Lr * H * sizeof(T), &bars[1]);
    }
    // thread mapping of phase (3): (row group g2, 16-byte chunk hc of the H axis)
    const int HV = H / VEC;
    const int G2 = p.G2;
    const int g2 = tid / HV, hc = tid - g2 * HV;
    const bool act3 = g2 < G2;
    float* dwa_img = p.dwa_acc + ((size_t)b * L + l0) * H;
    float4 old[RP][VEC / 4];
    if (act3 && !p.first) {
#pragma unroll
        for (int rr = 0; rr < RP; ++rr) {
            const int li = g2 + rr * G2;
            if (li < Lr) {
#pragma unroll
                for (int q = 0; q < VEC / 4; ++q)
                    old[rr][q] = *reinterpret_cast<const float4*>(dwa_img + (size_t)li * H + hc * VEC + 4 * q);
            }
        }
    }
    // (0) per-image vectors
    {
        const int d0 = min(D, rank * p.Ds), d1 = min(D, d0 + p.Ds);
        for (int d = tid; d < D; d += THREADS) {
            const float dzv = p.dz[(int64_t)b * p.ld_dz + d];
            float g = 1.0f;
            if (p.gate) g = sat_sigmoid<PRECISE>(p.uf[(int64_t)b * p.ld_uf + H + d]);
            float dc = dzv * g;
            if (p.dctx_extra) dc += p.dctx_extra[(int64_t)b * p.ld_dce + d];
            dctx_s[d] = from_f32<T>(dc);
            if (d >= d0 && d < d1) {
                float dfb = 0.0f;
                if (p.gate) dfb = dzv * to_f32<T>(reinterpret_cast<const T*>(p.ctx)[(int64_t)b * p.ld_ctx + d]) * g * (1.0f - g);
                reinterpret_cast<T*>(p.duf)[(int64_t)b * p.ld_duf + H + d] = from_f32<T>(dfb);
            }
        }
        for (int i = tid; i < H; i += THREADS) { v_s[i] = p.v[i]; uh_s[i] = p.uf[(int64_t)b * p.ld_uf + i]; }
        for (int i = tid; i < Lr; i += THREADS) al_s[i] = p.alpha[(int64_t)b * p.ld_alpha + l0 + i];
    }
    __syncthreads();
    mbar_wait(&bars[0], 0);

    // (1) dalpha_l = dctx . a_l (+ external grad); optional d a
    const int nchunks = D / VEC;
    float part = 0.0f;
    for (int li = warp; li < Lr; li += NWARPS) {
        float s = 0.0f;
        for (int ch = lane; ch < nchunks; ch += 32) {
            const uint4 u = *reinterpret_cast<const uint4*>(a_s + (size_t)li * D + ch * VEC);
            const uint4 w = *reinterpret_cast<const uint4*>(dctx_s + ch * VEC);
            float f[VEC], dc[VEC];
            Vec16<T>::unpack(u, f);
            Vec16<T>::unpack(w, dc);
#pragma unroll
            for (int i = 0; i < VEC; ++i) s = fmaf(f[i], dc[i], s);
        }
        s = warp_sum(s);
        if (p.galpha) s += p.galpha[(int64_t)b * p.ld_ga + l0 + li];
        const float al = al_s[li];
        if (lane == 0) de_s[li] = s;
        part += al * s;     // identical in all lanes
        if (p.da_acc) {
            float* dst = p.da_acc + ((size_t)b * L + l0 + li) * D;
            for (int d = lane; d < D; d += 32) {
                const float val = al * to_f32<T>(dctx_s[d]);
                dst[d] = p.first ? val : dst[d] + val;
            }
        }
    }
    if (lane == 0) misc_s[4 + warp] = part;
    __syncthreads();
    if (tid == 0) {
        float t = 0.0f;
        for (int w = 0; w < NWARPS; ++w) t += misc_s[4 + w];
        misc_s[0] = t;
    }
    cluster_sync_all();
    if (tid == 0) {
        float S = 0.0f;
        for (int r = 0; r < C; ++r) S += dsmem_ld_f32(dsmem_addr(misc_s, r));
        misc_s[2] = S;
    }
    __syncthreads();
    {
        const float S = misc_s[2];
        for (int i = tid; i < Lr; i += THREADS) de_s[i] = al_s[i] * (de_s[i] - S);   // softmax backward
    }
    __syncthreads();
    mbar_wait(&bars[1], 0);

    // (3) pass over staged Wa; thread (g2, hc) walks rows g2, g2+G2, ...
    float dU[VEC], dV[VEC];
#pragma unroll
    for (int i = 0; i < VEC; ++i) { dU[i] = 0.0f; dV[i] = 0.0f; }
    if (act3) {
        float uh[VEC], vv[VEC];
#pragma unroll
        for (int i = 0; i < VEC; ++i) { uh[i] = uh_s[hc * VEC + i]; vv[i] = v_s[hc * VEC + i]; }
        int rr = 0;
        for (int li = g2; li < Lr; li += G2, ++rr) {
            const float de = de_s[li];
            const uint4 u = *reinterpret_cast<const uint4*>(wa_s + (size_t)li * H + hc * VEC);
            float w[VEC], dp[VEC];
            Vec16<T>::unpack(u, w);
#pragma unroll
            for (int i = 0; i < VEC; ++i) {
                const float th = attn_tanh<PRECISE>(w[i] + uh[i]);
                dp[i] = de * vv[i] * (1.0f - th * th);
                dU[i] += dp[i];
                dV[i] = fmaf(de, th, dV[i]);
            }
            float* drow = dwa_img + (size_t)li * H + hc * VEC;
#pragma unroll
            for (int q = 0; q < VEC / 4; ++q) {
                float4 o = make_float4(dp[4 * q], dp[4 * q + 1], dp[4 * q + 2], dp[4 * q + 3]);
                if (!p.first) {
                    float4 prev;
                    if (rr < RP) {
                        prev = old[0][q];
#pragma unroll
                        for (int k = 1; k < RP; ++k) if (rr == k) prev = old[k][q];
                    } else {
                        prev = *reinterpret_cast<const float4*>(drow + 4 * q);
                    }
                    o.x += prev.x; o.y += prev.y; o.z += prev.z; o.w += prev.w;
                }
                *reinterpret_cast<float4*>(drow + 4 * q) = o;
            }
        }
    }
    __syncthreads();                 // a_s is dead from here on: red_s may alias it
    if (act3) {
#pragma unroll
        for (int i = 0; i < VEC; ++i) {
            red_s[((size_t)g2 * 2) * H + hc * VEC + i] = dU[i];
            red_s[((size_t)g2 * 2 + 1) * H + hc * VEC + i] = dV[i];
        }
    }
    __syncthreads();
    for (int i = tid; i < 2 * H; i += THREADS) {          // deterministic cross-group sum into group 0
        float t = red_s[i];
        for (int g = 1; g < G2; ++g) t += red_s[(size_t)g * 2 * H + i];
        red_s[i] = t;
    }
    if (tid == 0) {
        float t = 0.0f;
        for (int i = 0; i < Lr; ++i) t += de_s[i];
        misc_s[1] = t;
    }
    cluster_sync_all();
    {
        const int h0 = min(H, rank * p.Hs), h1 = min(H, h0 + p.Hs);
        for (int h = h0 + tid; h < h1; h += THREADS) {
            float du = 0.0f, dv = 0.0f;
#pragma unroll 4
            for (int r = 0; r < C; ++r) {
                du += dsmem_ld_f32(dsmem_addr(red_s + h, r));
                dv += dsmem_ld_f32(dsmem_addr(red_s + H + h, r));
            }
            reinterpret_cast<T*>(p.duf)[(int64_t)b * p.ld_duf + h] = from_f32<T>(du);
            float* dst = p.dv_acc + (size_t)b * H + h;
            *dst = p.first ? dv : *dst + dv;
        }
        if (rank == 0 && tid == 0) {
            float t = 0.0f;
            for (int r = 0; r < C; ++r) t += dsmem_ld_f32(dsmem_addr(misc_s + 1, r));
            p.dbv_acc[b] = p.first ? t : p.dbv_acc[b] + t;
        }
    }
    cluster_sync_all();
}

// ---- host launchers -------------------------------------------------------------------------------
// Cluster size: the smallest power of two whose per-CTA staged bytes fit `target` (so that several CTAs share
// an SM and their loads overlap each other's compute), keeping >= 2 locations per CTA.
inline void pick_cluster(int L, size_t row_bytes, size_t target, int maxc, int* C_out, int* Lc_out) {
    int C = 1;
    while (C < maxc) {
        const int Lc = (L + C - 1) / C;
        if ((size_t)Lc * row_bytes <= target) break;
        if ((L + 2 * C - 1) / (2 * C) < 2) break;
        C *= 2;
    }
    *C_out = C;
    *Lc_out = (L + C - 1) / C;
}

template <typename KernT>
inline int launch_cluster(KernT kern, dim3 grid, int C, size_t smem, cudaStream_t stream, void** args) {
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = grid;
    cfg.blockDim = dim3(THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = C; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    SAT_CHECK_CUDA(cudaLaunchKernelExC(&cfg, (const void*)kern, args));
    return SAT_OK;
}

// one-time per-kernel setup: opt in to 227 KB dynamic smem and 16-CTA clusters; returns the largest usable cluster
template <typename KernT>
inline int kernel_setup(KernT kern, int* maxc_out) {
    SAT_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    int maxc = 8;
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) == cudaSuccess) {
        // probe: can a 16-CTA cluster of this kernel be co-scheduled at a modest smem size?
        cudaLaunchConfig_t cfg;
        memset(&cfg, 0, sizeof(cfg));
        cfg.gridDim = dim3(16, 1); cfg.blockDim = dim3(THREADS); cfg.dynamicSmemBytes = 72 * 1024;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = 16; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        int nclusters = 0;
        if (cudaOccupancyMaxActiveClusters(&nclusters, (const void*)kern, &cfg) == cudaSuccess && nclusters >= 4) maxc = 16;
    }
    cudaGetLastError();
    *maxc_out = maxc;
    return SAT_OK;
}

constexpr size_t STAGE_TARGET = 68 * 1024;     // staged bytes per CTA -> 3 CTAs per SM

template <typename T, int NBG, bool PRECISE>
int launch_fwd_t(FwdParams p, cudaStream_t stream) {
    constexpr int VEC = Vec16<T>::N;
    auto kern = fwd_kernel<T, NBG, PRECISE>;
    static int maxc_dev[64];          // per device (kernel attributes belong to a device context)
    static SatPerDevice configured;
    int dev = 0;
    SAT_CHECK_CUDA(cudaGetDevice(&dev));
    SAT_TRY(configured.once([&]() -> int { return kernel_setup(kern, &maxc_dev[dev & 63]); }));
    const int maxc = maxc_dev[dev & 63];
    pick_cluster(p.L, (size_t)(p.D + p.H) * sizeof(T), STAGE_TARGET, maxc, &p.C, &p.Lc);
    const int nchunks = p.D / VEC;
    p.G = 1;
    if (nchunks < THREADS) { p.G = THREADS / nchunks; if (p.G > GMAX) p.G = GMAX; if (p.G < 1) p.G = 1; }
    p.Ds = (((p.D + p.C - 1) / p.C) + 3) & ~3;
    SAT_REQUIRE(nchunks <= 2 * THREADS, "attention fwd: D=%d too wide for this build", p.D);
    SmemLayout sl = fwd_layout<T>(NBG, p.NB, p.Lc, p.D, p.H, p.G);
    SAT_REQUIRE(sl.total <= 227 * 1024, "attention fwd: staged chunk needs %zu B shared memory (> 227 KB): L=%d D=%d", sl.total, p.L, p.D);
    void* args[] = {&p, &sl};
    SatProfScope prof(SAT_PROF_ATTN_FWD, stream);
    return launch_cluster(kern, dim3(p.C, p.NI), p.C, sl.total, stream, args);
}

template <typename T, bool PRECISE>
int launch_fwd(FwdParams p, cudaStream_t stream) {
    SAT_REQUIRE(p.D % Vec16<T>::N == 0 && p.H % Vec16<T>::N == 0, "attention: D and H must be multiples of %d", Vec16<T>::N);
    SAT_REQUIRE(p.NB >= 1, "attention: NB >= 1");
    if (p.NB == 1) return launch_fwd_t<T, 1, PRECISE>(p, stream);
    if (p.NB <= 3) return launch_fwd_t<T, 3, PRECISE>(p, stream);
    return launch_fwd_t<T, 5, PRECISE>(p, stream);
}

template <typename T, bool PRECISE>
int launch_bwd(BwdParams p, cudaStream_t stream) {
    constexpr int VEC = Vec16<T>::N;
    SAT_REQUIRE(p.D % VEC == 0 && p.H % VEC == 0, "attention: D and H must be multiples of %d", VEC);
    SAT_REQUIRE(p.H / VEC <= THREADS, "attention bwd: H=%d too wide for this build", p.H);
    auto kern = bwd_kernel<T, PRECISE>;
    static int maxc_dev[64];          // per device (kernel attributes belong to a device context)
    static SatPerDevice configured;
    int dev = 0;
    SAT_CHECK_CUDA(cudaGetDevice(&dev));
    SAT_TRY(configured.once([&]() -> int { return kernel_setup(kern, &maxc_dev[dev & 63]); }));
    const int maxc = maxc_dev[dev & 63];
    pick_cluster(p.L, (size_t)(p.D + p.H) * sizeof(T), STAGE_TARGET, maxc, &p.C, &p.Lc);
    p.Ds = (p.D + p.C - 1) / p.C;
    p.Hs = (p.H + p.C - 1) / p.C;
    p.G2 = THREADS / (p.H / VEC);
    if (p.G2 > GMAX) p.G2 = GMAX;
    BwdSmem sl = bwd_layout<T>(p.Lc, p.D, p.H, p.G2);
    SAT_REQUIRE(sl.total <= 227 * 1024, "attention bwd: needs %zu B shared memory (> 227 KB)", sl.total);
    void* args[] = {&p, &sl};
    SatProfScope prof(SAT_PROF_ATTN_BWD, stream);
    return launch_cluster(kern, dim3(p.C, p.B), p.C, sl.total, stream, args);
}
}  // namespace attn
