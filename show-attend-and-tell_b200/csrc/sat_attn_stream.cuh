// Streaming (persistent, warp-specialised) variants of the fused attention decode step -- the fast path for the
// shapes the configs use (D in {512, 2048}, H in {512, 768}).  sat_attn.cuh keeps the generic cluster kernels.
//
// One persistent CTA per SM walks a static round-robin list of work items (item = one image, or one image x one
// group of beam rows).  Warp 8 is the PRODUCER: a single elected lane keeps a ring of shared-memory slots full
// with cp.async.bulk (TMA engine, mbarrier complete_tx) -- it runs ahead across item boundaries, so HBM requests
// never stop while the consumers reduce / finalise an item.  Warps 0-7 are CONSUMERS: slot n of the global slot
// sequence belongs to warp n % 8, which owns a private online-softmax state (running max m, running sum s and an
// un-normalised context accumulator of D/32 fp32 registers per lane), so no block-wide synchronisation happens
// while an image streams; per item the eight partial states are merged through shared memory (3 named barriers).
// Every element of `a` and `Wa` crosses HBM->SM exactly once per step: L*(H+D)*sizeof(T) bytes per image.
#pragma once
#include "sat_attn.cuh"

extern int g_attn_generic;     // debug knob (sat_set_attn_generic): 1 = always use the generic cluster kernels

namespace attn_s {
using attn::FwdParams;
using attn::BwdParams;

constexpr int CWARPS = 8;
constexpr int CTHREADS = CWARPS * 32;
constexpr int THREADS = CTHREADS + 32;
constexpr int MAX_SLOTS = 32;
constexpr size_t SMEM_MAX = 227 * 1024;

__device__ __forceinline__ void consumer_bar() { asm volatile("bar.sync 1, %0;" ::"n"(CTHREADS) : "memory"); }

constexpr int pow2floor(int x) { int r = 1; while (2 * r <= x) r *= 2; return r; }
constexpr int clampi(int x, int lo, int hi) { return x < lo ? lo : (x > hi ? hi : x); }

template <typename T, int D, int H> struct Geo {
    static constexpr int VEC = Vec16<T>::N;
    static constexpr int DPL = D / VEC / 32;      // 16-byte chunks per lane per feature row
    static constexpr int HPL = H / VEC / 32;      // 16-byte chunks per lane per projection row
    static constexpr int ROWB = (H + D) * (int)sizeof(T);
    static constexpr int LS = clampi(pow2floor(12288 / ROWB), 2, 8);      // locations per forward slot
    static constexpr int FWD_SLOT = LS * ROWB;
    // backward: phase A streams feature rows, phase B streams projection rows, both through equal-sized slots
    static constexpr int BWD_SLOT = (2 * D * (int)sizeof(T) > 8192) ? 2 * D * (int)sizeof(T) : 8192;
    static constexpr int LSA = BWD_SLOT / (D * (int)sizeof(T));
    static constexpr int LSB = clampi(pow2floor(BWD_SLOT / (H * (int)sizeof(T))), 2, 8);
    static_assert(D % (VEC * 32) == 0 && H % (VEC * 32) == 0, "D and H must be multiples of 32 16-byte chunks");
};

template <typename T> __device__ __forceinline__ uint4 lds16(const T* p) { return *reinterpret_cast<const uint4*>(p); }

// ------------------------------------------------------------------------------------------------------------
// forward
// ------------------------------------------------------------------------------------------------------------
struct FwdSmem { size_t ring, red, es, wms, bars, total; int nslots, lpad; };

template <typename T, int D, int H, int NBG>
inline FwdSmem fwd_smem(int L) {
    using G = Geo<T, D, H>;
    FwdSmem s;
    s.lpad = (L + 3) & ~3;
    const size_t red = (size_t)4 * NBG * D * 4, es = (size_t)2 * NBG * s.lpad * 4, wms = (size_t)2 * CWARPS * NBG * 2 * 4;
    const size_t fixed = red + es + wms + 2 * MAX_SLOTS * 8 + 256;
    int ns = (int)((SMEM_MAX - fixed) / G::FWD_SLOT);
    ns = (ns / CWARPS) * CWARPS;
    if (ns > MAX_SLOTS) ns = MAX_SLOTS;
    s.nslots = ns;
    size_t o = 0;
    s.ring = o; o += (size_t)(ns > 0 ? ns : 0) * G::FWD_SLOT;
    o = (o + 127) & ~(size_t)127;
    s.red = o; o += red;
    s.es = o; o += es;
    s.wms = o; o += wms;
    o = (o + 15) & ~(size_t)15;
    s.bars = o; o += 2 * MAX_SLOTS * 8;
    s.total = o;
    return s;
}

template <typename T, int D, int H, int NBG, bool PRECISE>
__global__ void __launch_bounds__(THREADS, 1)
fwd_stream_kernel(FwdParams p, FwdSmem sl, int nitems, int ngroups) {
    using G = Geo<T, D, H>;
    constexpr int VEC = G::VEC, DPL = G::DPL, HPL = G::HPL, LS = G::LS;
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t* ring = smem + sl.ring;
    float* red = reinterpret_cast<float*>(smem + sl.red);          // [4][NBG][D]
    float* e_s = reinterpret_cast<float*>(smem + sl.es);           // [2][NBG][lpad]
    float* wms = reinterpret_cast<float*>(smem + sl.wms);          // [2][CWARPS][NBG][2]
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + sl.bars);
    uint64_t* empty = full + MAX_SLOTS;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int L = p.L, nslots = sl.nslots, lpad = sl.lpad;
    const int spi = (L + LS - 1) / LS;                              // slots per item

    if (tid == 0) {
        for (int i = 0; i < nslots; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
        mbar_fence_init();
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();

    if (warp == CWARPS) {
        // ---------------- producer ----------------
        if (lane == 0) {
            uint32_t n = 0;
            for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
                const int img = item / ngroups;
                const T* wa_g = reinterpret_cast<const T*>(p.wa) + (size_t)img * L * H;
                const T* a_g = reinterpret_cast<const T*>(p.a) + (size_t)img * L * D;
                for (int k = 0; k < spi; ++k, ++n) {
                    const uint32_t slot = n % nslots, ph = (n / nslots) & 1;
                    mbar_wait(&empty[slot], ph ^ 1);
                    const int l0 = k * LS, nloc = min(LS, L - l0);
                    uint8_t* dst = ring + (size_t)slot * G::FWD_SLOT;
                    mbar_expect_tx(&full[slot], (uint32_t)(nloc * G::ROWB));
                    bulk_g2s(dst, wa_g + (size_t)l0 * H, (uint32_t)(nloc * H * sizeof(T)), &full[slot]);
                    bulk_g2s(dst + LS * H * sizeof(T), a_g + (size_t)l0 * D, (uint32_t)(nloc * D * sizeof(T)), &full[slot]);
                }
            }
        }
        return;
    }

    // ---------------- consumers ----------------
    float vreg[HPL * VEC];
#pragma unroll
    for (int j = 0; j < HPL; ++j)
#pragma unroll
        for (int q = 0; q < VEC / 4; ++q) {
            const float4 t = *reinterpret_cast<const float4*>(p.v + (lane + 32 * j) * VEC + 4 * q);
            vreg[j * VEC + 4 * q] = t.x; vreg[j * VEC + 4 * q + 1] = t.y; vreg[j * VEC + 4 * q + 2] = t.z; vreg[j * VEC + 4 * q + 3] = t.w;
        }
    const float bv = p.bv[0];
    uint32_t n_base = 0;
    int par = 0;
    for (int item = blockIdx.x; item < nitems; item += gridDim.x, n_base += spi, par ^= 1) {
        const int img = item / ngroups, grp = item - img * ngroups;
        const int row_base = img * p.NB + grp * NBG;
        const int nb_valid = min(NBG, p.NB - grp * NBG);
        float uh[HPL * VEC];
        if (NBG == 1) {
#pragma unroll
            for (int j = 0; j < HPL; ++j)
#pragma unroll
                for (int q = 0; q < VEC / 4; ++q) {
                    const float4 t = *reinterpret_cast<const float4*>(p.uf + (int64_t)row_base * p.ld_uf + (lane + 32 * j) * VEC + 4 * q);
                    uh[j * VEC + 4 * q] = t.x; uh[j * VEC + 4 * q + 1] = t.y; uh[j * VEC + 4 * q + 2] = t.z; uh[j * VEC + 4 * q + 3] = t.w;
                }
        }
        float m[NBG], s[NBG], acc[NBG][DPL * VEC];
#pragma unroll
        for (int nb = 0; nb < NBG; ++nb) {
            m[nb] = -INFINITY; s[nb] = 0.0f;
#pragma unroll
            for (int i = 0; i < DPL * VEC; ++i) acc[nb][i] = 0.0f;
        }
        float* e_item = e_s + (size_t)par * NBG * lpad;

        const int first_k = (int)((warp + CWARPS - (n_base % CWARPS)) % CWARPS);
        for (int k = first_k; k < spi; k += CWARPS) {
            const uint32_t n = n_base + k, slot = n % nslots, ph = (n / nslots) & 1;
            mbar_wait(&full[slot], ph);
            const T* wa_s = reinterpret_cast<const T*>(ring + (size_t)slot * G::FWD_SLOT);
            const T* a_s = wa_s + LS * H;
            const int l0 = k * LS, nloc = min(LS, L - l0);
#pragma unroll 1
            for (int lp = 0; lp < nloc; lp += 2) {
                const bool two = lp + 1 < nloc;
                const int lq = two ? lp + 1 : lp;
                float sc0[NBG], sc1[NBG];
#pragma unroll
                for (int nb = 0; nb < NBG; ++nb) { sc0[nb] = 0.0f; sc1[nb] = 0.0f; }
#pragma unroll
                for (int j = 0; j < HPL; ++j) {
                    const int c = lane + 32 * j;
                    float f0[VEC], f1[VEC];
                    Vec16<T>::unpack(lds16(wa_s + (size_t)lp * H + c * VEC), f0);
                    Vec16<T>::unpack(lds16(wa_s + (size_t)lq * H + c * VEC), f1);
#pragma unroll
                    for (int nb = 0; nb < NBG; ++nb) {
                        float u[VEC];
                        if (NBG == 1) {
#pragma unroll
                            for (int i = 0; i < VEC; ++i) u[i] = uh[j * VEC + i];
                        } else {
                            const int64_t row = row_base + min(nb, nb_valid - 1);
#pragma unroll
                            for (int q = 0; q < VEC / 4; ++q) {
                                const float4 t = __ldg(reinterpret_cast<const float4*>(p.uf + row * p.ld_uf + c * VEC + 4 * q));
                                u[4 * q] = t.x; u[4 * q + 1] = t.y; u[4 * q + 2] = t.z; u[4 * q + 3] = t.w;
                            }
                        }
#pragma unroll
                        for (int i = 0; i < VEC; ++i) {
                            sc0[nb] = fmaf(vreg[j * VEC + i], sat_tanh<PRECISE>(f0[i] + u[i]), sc0[nb]);
                            sc1[nb] = fmaf(vreg[j * VEC + i], sat_tanh<PRECISE>(f1[i] + u[i]), sc1[nb]);
                        }
                    }
                }
                float p0[NBG], p1[NBG];
#pragma unroll
                for (int nb = 0; nb < NBG; ++nb) {
                    const float e0 = warp_sum(sc0[nb]) + bv, e1 = warp_sum(sc1[nb]) + bv;
                    if (lane == 0) { e_item[nb * lpad + l0 + lp] = e0; if (two) e_item[nb * lpad + l0 + lp + 1] = e1; }
                    const float mn = fmaxf(m[nb], two ? fmaxf(e0, e1) : e0);
                    if (mn != m[nb]) {                                    // warp-uniform: rescale the running state
                        const float r = sat_exp<PRECISE>(m[nb] - mn);
                        s[nb] *= r;
#pragma unroll
                        for (int i = 0; i < DPL * VEC; ++i) acc[nb][i] *= r;
                        m[nb] = mn;
                    }
                    p0[nb] = sat_exp<PRECISE>(e0 - mn);
                    p1[nb] = two ? sat_exp<PRECISE>(e1 - mn) : 0.0f;
                    s[nb] += p0[nb] + p1[nb];
                }
#pragma unroll
                for (int j = 0; j < DPL; ++j) {
                    const int c = lane + 32 * j;
                    float g0[VEC], g1[VEC];
                    Vec16<T>::unpack(lds16(a_s + (size_t)lp * D + c * VEC), g0);
                    Vec16<T>::unpack(lds16(a_s + (size_t)lq * D + c * VEC), g1);
#pragma unroll
                    for (int nb = 0; nb < NBG; ++nb)
#pragma unroll
                        for (int i = 0; i < VEC; ++i)
                            acc[nb][j * VEC + i] = fmaf(p0[nb], g0[i], fmaf(p1[nb], g1[i], acc[nb][j * VEC + i]));
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[slot]);
        }

        // ---- merge the eight per-warp online-softmax states of this item ----
        float* wms_item = wms + (size_t)par * CWARPS * NBG * 2;
        if (lane == 0) {
#pragma unroll
            for (int nb = 0; nb < NBG; ++nb) { wms_item[(warp * NBG + nb) * 2] = m[nb]; wms_item[(warp * NBG + nb) * 2 + 1] = s[nb]; }
        }
        consumer_bar();                                                   // B1: all (m, s) and all scores are published
        float Mx[NBG], inv[NBG];
#pragma unroll
        for (int nb = 0; nb < NBG; ++nb) {
            float M = -INFINITY;
#pragma unroll
            for (int w = 0; w < CWARPS; ++w) M = fmaxf(M, wms_item[(w * NBG + nb) * 2]);
            float S = 0.0f;
#pragma unroll
            for (int w = 0; w < CWARPS; ++w) S = fmaf(wms_item[(w * NBG + nb) * 2 + 1], sat_exp<PRECISE>(wms_item[(w * NBG + nb) * 2] - M), S);
            Mx[nb] = M; inv[nb] = 1.0f / S;
            const float f = sat_exp<PRECISE>(m[nb] - M);
#pragma unroll
            for (int i = 0; i < DPL * VEC; ++i) acc[nb][i] *= f;
        }
        for (int i = tid; i < nb_valid * L; i += CTHREADS) {              // alphas (fp32, decoder.py:128)
            const int nb = i / L, l = i - nb * L;
            float M = Mx[0], iv = inv[0];
#pragma unroll
            for (int k = 1; k < NBG; ++k) if (nb == k) { M = Mx[k]; iv = inv[k]; }
            p.alpha[(int64_t)(row_base + nb) * p.ld_alpha + l] = sat_exp<PRECISE>(e_item[nb * lpad + l] - M) * iv;
        }
        auto red_ptr = [&](int w, int nb, int j, int q) { return red + ((size_t)(w * NBG + nb) * D + (lane + 32 * j) * VEC + 4 * q); };
        if (warp >= 4) {
#pragma unroll
            for (int nb = 0; nb < NBG; ++nb)
#pragma unroll
                for (int j = 0; j < DPL; ++j)
#pragma unroll
                    for (int q = 0; q < VEC / 4; ++q)
                        *reinterpret_cast<float4*>(red_ptr(warp - 4, nb, j, q)) =
                            make_float4(acc[nb][j * VEC + 4 * q], acc[nb][j * VEC + 4 * q + 1], acc[nb][j * VEC + 4 * q + 2], acc[nb][j * VEC + 4 * q + 3]);
        }
        consumer_bar();                                                   // B2
        if (warp < 4) {
#pragma unroll
            for (int nb = 0; nb < NBG; ++nb)
#pragma unroll
                for (int j = 0; j < DPL; ++j)
#pragma unroll
                    for (int q = 0; q < VEC / 4; ++q) {
                        float4* rp = reinterpret_cast<float4*>(red_ptr(warp, nb, j, q));
                        const float4 o = *rp;
                        *rp = make_float4(acc[nb][j * VEC + 4 * q] + o.x, acc[nb][j * VEC + 4 * q + 1] + o.y,
                                          acc[nb][j * VEC + 4 * q + 2] + o.z, acc[nb][j * VEC + 4 * q + 3] + o.w);
                    }
        }
        consumer_bar();                                                   // B3
        for (int i = tid; i < nb_valid * (D / 4); i += CTHREADS) {
            const int nb = i / (D / 4), d = 4 * (i - nb * (D / 4));
            float iv = inv[0];
#pragma unroll
            for (int k = 1; k < NBG; ++k) if (nb == k) iv = inv[k];
            float4 t = *reinterpret_cast<const float4*>(red + (size_t)nb * D + d);
#pragma unroll
            for (int w = 1; w < 4; ++w) {
                const float4 x = *reinterpret_cast<const float4*>(red + (size_t)(w * NBG + nb) * D + d);
                t.x += x.x; t.y += x.y; t.z += x.z; t.w += x.w;
            }
            const int64_t row = row_base + nb;
            float c4[4] = {t.x * iv, t.y * iv, t.z * iv, t.w * iv};
            if (p.ctx) store4<T>(reinterpret_cast<T*>(p.ctx) + row * p.ld_ctx + d, c4);
            if (p.z) {
                if (p.gate) {
                    const float4 gp = *reinterpret_cast<const float4*>(p.uf + row * p.ld_uf + H + d);
                    c4[0] *= sat_sigmoid<PRECISE>(gp.x); c4[1] *= sat_sigmoid<PRECISE>(gp.y);
                    c4[2] *= sat_sigmoid<PRECISE>(gp.z); c4[3] *= sat_sigmoid<PRECISE>(gp.w);
                }
                store4<T>(reinterpret_cast<T*>(p.z) + row * p.ld_z + d, c4);
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------------------
// backward of one attention step (NB == 1), same producer/consumer structure.  Per item (image b):
//   phase A  stream the feature rows a_l:  dalpha_l = dctx . a_l (+ external grad), per-warp sum_l alpha_l dalpha_l
//   barrier  S = sum over warps          (softmax backward needs the full sum before any de_l)
//   phase B  stream the projection rows Wa_l:  de_l = alpha_l (dalpha_l - S);  th = tanh(Wa_l + Uh);
//            dpre_l = de_l v (1 - th^2)  -> written ONCE, in T, to this step's slice of the (T,B,L,H) dpre buffer
//            (the time sum is taken by one pass at the end of BPTT instead of a fp32 read-modify-write per step);
//            per-warp dUh += dpre_l, dv += de_l th, sum de_l
//   barrier  cross-warp reduction of dUh / dv through shared memory.
// HBM traffic per image per step: L (H+D) sizeof(T) read + L H sizeof(T) written.
// ------------------------------------------------------------------------------------------------------------
struct BwdSmemS { size_t ring, red, de, wsum, bars, total; int nslots, lpad; };

template <typename T, int D, int H>
inline BwdSmemS bwd_smem(int L) {
    using G = Geo<T, D, H>;
    BwdSmemS s;
    s.lpad = (L + 3) & ~3;
    const size_t red = (size_t)CWARPS * 2 * H * 4, de = (size_t)2 * s.lpad * 4, wsum = (size_t)2 * CWARPS * 2 * 4;
    const size_t fixed = red + de + wsum + 2 * MAX_SLOTS * 8 + 256;
    int ns = (int)((SMEM_MAX - fixed) / G::BWD_SLOT);
    ns = (ns / CWARPS) * CWARPS;
    if (ns > MAX_SLOTS) ns = MAX_SLOTS;
    s.nslots = ns;
    size_t o = 0;
    s.ring = o; o += (size_t)(ns > 0 ? ns : 0) * G::BWD_SLOT;
    o = (o + 127) & ~(size_t)127;
    s.red = o; o += red;
    s.de = o; o += de;
    s.wsum = o; o += wsum;
    o = (o + 15) & ~(size_t)15;
    s.bars = o; o += 2 * MAX_SLOTS * 8;
    s.total = o;
    return s;
}

template <typename T, int D, int H, bool PRECISE>
__global__ void __launch_bounds__(THREADS, 1)
bwd_stream_kernel(BwdParams p, BwdSmemS sl) {
    using G = Geo<T, D, H>;
    constexpr int VEC = G::VEC, DPL = G::DPL, HPL = G::HPL, LSA = G::LSA, LSB = G::LSB;
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t* ring = smem + sl.ring;
    float* red = reinterpret_cast<float*>(smem + sl.red);          // [CWARPS][2][H]
    float* de_s = reinterpret_cast<float*>(smem + sl.de);          // [2][lpad]
    float* wsum = reinterpret_cast<float*>(smem + sl.wsum);        // [2][CWARPS][2]
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + sl.bars);
    uint64_t* empty = full + MAX_SLOTS;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int L = p.L, nslots = sl.nslots, lpad = sl.lpad, nitems = p.B;
    const int spa = (L + LSA - 1) / LSA, spb = (L + LSB - 1) / LSB, spi = spa + spb;

    if (tid == 0) {
        for (int i = 0; i < nslots; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
        mbar_fence_init();
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();

    if (warp == CWARPS) {
        if (lane == 0) {
            uint32_t n = 0;
            for (int b = blockIdx.x; b < nitems; b += gridDim.x) {
                const T* a_g = reinterpret_cast<const T*>(p.a) + (size_t)b * L * D;
                const T* wa_g = reinterpret_cast<const T*>(p.wa) + (size_t)b * L * H;
                for (int k = 0; k < spi; ++k, ++n) {
                    const uint32_t slot = n % nslots, ph = (n / nslots) & 1;
                    mbar_wait(&empty[slot], ph ^ 1);
                    uint8_t* dst = ring + (size_t)slot * G::BWD_SLOT;
                    if (k < spa) {
                        const int l0 = k * LSA, nloc = min(LSA, L - l0);
                        const uint32_t bytes = (uint32_t)(nloc * D * sizeof(T));
                        mbar_expect_tx(&full[slot], bytes);
                        bulk_g2s(dst, a_g + (size_t)l0 * D, bytes, &full[slot]);
                    } else {
                        const int l0 = (k - spa) * LSB, nloc = min(LSB, L - l0);
                        const uint32_t bytes = (uint32_t)(nloc * H * sizeof(T));
                        mbar_expect_tx(&full[slot], bytes);
                        bulk_g2s(dst, wa_g + (size_t)l0 * H, bytes, &full[slot]);
                    }
                }
            }
        }
        return;
    }

    float vreg[HPL * VEC];
#pragma unroll
    for (int j = 0; j < HPL; ++j)
#pragma unroll
        for (int q = 0; q < VEC / 4; ++q) {
            const float4 t = *reinterpret_cast<const float4*>(p.v + (lane + 32 * j) * VEC + 4 * q);
            vreg[j * VEC + 4 * q] = t.x; vreg[j * VEC + 4 * q + 1] = t.y; vreg[j * VEC + 4 * q + 2] = t.z; vreg[j * VEC + 4 * q + 3] = t.w;
        }
    uint32_t n_base = 0;
    int par = 0;
    for (int b = blockIdx.x; b < nitems; b += gridDim.x, n_base += spi, par ^= 1) {
        float* de_item = de_s + (size_t)par * lpad;
        float* ws_item = wsum + (size_t)par * CWARPS * 2;
        const float* alpha_g = p.alpha + (int64_t)b * p.ld_alpha;
        float part = 0.0f;
        {
            // ---- phase A ----
            float dctx[DPL * VEC];
#pragma unroll
            for (int j = 0; j < DPL; ++j) {
                const int c = lane + 32 * j;
#pragma unroll
                for (int q = 0; q < VEC / 4; ++q) {
                    const int d = c * VEC + 4 * q;
                    const float4 dz4 = *reinterpret_cast<const float4*>(p.dz + (int64_t)b * p.ld_dz + d);
                    float dz[4] = {dz4.x, dz4.y, dz4.z, dz4.w}, g[4] = {1.f, 1.f, 1.f, 1.f};
                    if (p.gate) {
                        const float4 gp = *reinterpret_cast<const float4*>(p.uf + (int64_t)b * p.ld_uf + H + d);
                        g[0] = sat_sigmoid<PRECISE>(gp.x); g[1] = sat_sigmoid<PRECISE>(gp.y);
                        g[2] = sat_sigmoid<PRECISE>(gp.z); g[3] = sat_sigmoid<PRECISE>(gp.w);
                    }
                    float ex[4] = {0.f, 0.f, 0.f, 0.f};
                    if (p.dctx_extra) {
                        const float4 e4 = *reinterpret_cast<const float4*>(p.dctx_extra + (int64_t)b * p.ld_dce + d);
                        ex[0] = e4.x; ex[1] = e4.y; ex[2] = e4.z; ex[3] = e4.w;
                    }
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        dctx[j * VEC + 4 * q + i] = dz[i] * g[i] + ex[i];
                    }
                    if ((j % CWARPS) == warp) {                      // d f_beta pre-activation: written once per image
                        float dfb[4] = {0.f, 0.f, 0.f, 0.f};
                        if (p.gate) {
                            float cx[4];
                            load4<T>(reinterpret_cast<const T*>(p.ctx) + (int64_t)b * p.ld_ctx + d, cx);
#pragma unroll
                            for (int i = 0; i < 4; ++i) dfb[i] = dz[i] * cx[i] * g[i] * (1.0f - g[i]);
                        }
                        store4<T>(reinterpret_cast<T*>(p.duf) + (int64_t)b * p.ld_duf + H + d, dfb);
                    }
                }
            }
            const int first_k = (int)((warp + CWARPS - (n_base % CWARPS)) % CWARPS);
            for (int k = first_k; k < spa; k += CWARPS) {
                const uint32_t n = n_base + k, slot = n % nslots, ph = (n / nslots) & 1;
                mbar_wait(&full[slot], ph);
                const T* a_s = reinterpret_cast<const T*>(ring + (size_t)slot * G::BWD_SLOT);
                const int l0 = k * LSA, nloc = min(LSA, L - l0);
#pragma unroll 1
                for (int lp = 0; lp < nloc; lp += 2) {
                    const bool two = lp + 1 < nloc;
                    const int lq = two ? lp + 1 : lp;
                    float s0 = 0.0f, s1 = 0.0f;
#pragma unroll
                    for (int j = 0; j < DPL; ++j) {
                        const int c = lane + 32 * j;
                        float g0[VEC], g1[VEC];
                        Vec16<T>::unpack(lds16(a_s + (size_t)lp * D + c * VEC), g0);
                        Vec16<T>::unpack(lds16(a_s + (size_t)lq * D + c * VEC), g1);
#pragma unroll
                        for (int i = 0; i < VEC; ++i) { s0 = fmaf(g0[i], dctx[j * VEC + i], s0); s1 = fmaf(g1[i], dctx[j * VEC + i], s1); }
                    }
                    s0 = warp_sum(s0); s1 = warp_sum(s1);
                    const int l = l0 + lp;
                    if (p.galpha) { s0 += p.galpha[(int64_t)b * p.ld_ga + l]; if (two) s1 += p.galpha[(int64_t)b * p.ld_ga + l + 1]; }
                    part = fmaf(alpha_g[l], s0, part);
                    if (two) part = fmaf(alpha_g[l + 1], s1, part);
                    if (lane == 0) { de_item[l] = s0; if (two) de_item[l + 1] = s1; }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&empty[slot]);
            }
        }
        if (lane == 0) ws_item[warp * 2] = part;
        consumer_bar();                                                   // BA: all dalpha_l and partial sums published
        float S = 0.0f;
#pragma unroll
        for (int w = 0; w < CWARPS; ++w) S += ws_item[w * 2];

        // ---- phase B ----
        float uh[HPL * VEC], dU[HPL * VEC], dV[HPL * VEC];
#pragma unroll
        for (int j = 0; j < HPL; ++j)
#pragma unroll
            for (int q = 0; q < VEC / 4; ++q) {
                const float4 t = *reinterpret_cast<const float4*>(p.uf + (int64_t)b * p.ld_uf + (lane + 32 * j) * VEC + 4 * q);
                uh[j * VEC + 4 * q] = t.x; uh[j * VEC + 4 * q + 1] = t.y; uh[j * VEC + 4 * q + 2] = t.z; uh[j * VEC + 4 * q + 3] = t.w;
            }
#pragma unroll
        for (int i = 0; i < HPL * VEC; ++i) { dU[i] = 0.0f; dV[i] = 0.0f; }
        float desum = 0.0f;
        T* dpre_g = reinterpret_cast<T*>(p.dpre_out) + (size_t)b * L * H;
        {
            const uint32_t nb0 = n_base + spa;
            const int first_k = (int)((warp + CWARPS - (nb0 % CWARPS)) % CWARPS);
            for (int k = first_k; k < spb; k += CWARPS) {
                const uint32_t n = nb0 + k, slot = n % nslots, ph = (n / nslots) & 1;
                mbar_wait(&full[slot], ph);
                const T* wa_s = reinterpret_cast<const T*>(ring + (size_t)slot * G::BWD_SLOT);
                const int l0 = k * LSB, nloc = min(LSB, L - l0);
#pragma unroll 2
                for (int lp = 0; lp < nloc; ++lp) {
                    const int l = l0 + lp;
                    const float de = alpha_g[l] * (de_item[l] - S);          // softmax backward
                    desum += de;
#pragma unroll
                    for (int j = 0; j < HPL; ++j) {
                        const int c = lane + 32 * j;
                        float w[VEC], dp[VEC];
                        Vec16<T>::unpack(lds16(wa_s + (size_t)lp * H + c * VEC), w);
#pragma unroll
                        for (int i = 0; i < VEC; ++i) {
                            const float th = sat_tanh<PRECISE>(w[i] + uh[j * VEC + i]);
                            dp[i] = de * vreg[j * VEC + i] * (1.0f - th * th);
                            dU[j * VEC + i] += dp[i];
                            dV[j * VEC + i] = fmaf(de, th, dV[j * VEC + i]);
                        }
                        *reinterpret_cast<uint4*>(dpre_g + (size_t)l * H + c * VEC) = Vec16<T>::pack(dp);
                    }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&empty[slot]);
            }
        }
        // ---- cross-warp reduction of dUh / dv ----
#pragma unroll
        for (int j = 0; j < HPL; ++j)
#pragma unroll
            for (int q = 0; q < VEC / 4; ++q) {
                const int h = (lane + 32 * j) * VEC + 4 * q;
                *reinterpret_cast<float4*>(red + (size_t)(warp * 2) * H + h) =
                    make_float4(dU[j * VEC + 4 * q], dU[j * VEC + 4 * q + 1], dU[j * VEC + 4 * q + 2], dU[j * VEC + 4 * q + 3]);
                *reinterpret_cast<float4*>(red + (size_t)(warp * 2 + 1) * H + h) =
                    make_float4(dV[j * VEC + 4 * q], dV[j * VEC + 4 * q + 1], dV[j * VEC + 4 * q + 2], dV[j * VEC + 4 * q + 3]);
            }
        if (lane == 0) ws_item[warp * 2 + 1] = desum;
        consumer_bar();                                                   // BB
        for (int h = tid; h < H; h += CTHREADS) {
            float du = 0.0f, dv = 0.0f;
#pragma unroll
            for (int w = 0; w < CWARPS; ++w) { du += red[(size_t)(w * 2) * H + h]; dv += red[(size_t)(w * 2 + 1) * H + h]; }
            reinterpret_cast<T*>(p.duf)[(int64_t)b * p.ld_duf + h] = from_f32<T>(du);
            float* dst = p.dv_acc + (size_t)b * H + h;
            *dst = p.first ? dv : *dst + dv;
        }
        if (tid == 0) {
            float t = 0.0f;
#pragma unroll
            for (int w = 0; w < CWARPS; ++w) t += ws_item[w * 2 + 1];
            p.dbv_acc[b] = p.first ? t : p.dbv_acc[b] + t;
        }
    }
}

// ---- host side ------------------------------------------------------------------------------------------
inline int num_sms() {
    static int n = 0;
    if (!n) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) n = 148;
    }
    return n;
}
static inline bool aligned(const void* p, size_t a) { return ((uintptr_t)p % a) == 0; }
constexpr int MIN_ITEMS = 64;      // below this the generic cluster kernels spread one image over several SMs

inline bool shape_ok(int L, int D, int H) { return !g_attn_generic && (D == 512 || D == 2048) && (H == 512 || H == 768) && L >= 1 && L <= 2048; }

template <typename T>
inline bool fwd_eligible(const FwdParams& p) {
    if (!shape_ok(p.L, p.D, p.H)) return false;
    if (p.NB > 1 && p.D != 512) return false;
    const int ngroups = p.NB == 1 ? 1 : (p.NB + 4) / 5;
    if (p.NI * ngroups < MIN_ITEMS) return false;
    const size_t oa = sizeof(T) == 2 ? 8 : 16;
    return aligned(p.a, 16) && aligned(p.wa, 16) && aligned(p.uf, 16) && (p.ld_uf % 4 == 0) && aligned(p.v, 16) &&
           (!p.ctx || (aligned(p.ctx, oa) && p.ld_ctx % 4 == 0)) && (!p.z || (aligned(p.z, oa) && p.ld_z % 4 == 0));
}

template <typename T, int D, int H, int NBG, bool PRECISE>
int launch_fwd_inst(const FwdParams& p, cudaStream_t stream) {
    auto kern = fwd_stream_kernel<T, D, H, NBG, PRECISE>;
    static bool configured = false;
    if (!configured) {
        SAT_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_MAX));
        configured = true;
    }
    const FwdSmem sl = fwd_smem<T, D, H, NBG>(p.L);
    SAT_REQUIRE(sl.nslots >= CWARPS && sl.total <= SMEM_MAX, "attention fwd (streaming): shared memory budget exceeded (L=%d)", p.L);
    const int ngroups = (p.NB + NBG - 1) / NBG, nitems = p.NI * ngroups;
    const int grid = nitems < num_sms() ? nitems : num_sms();
    SatProfScope prof(SAT_PROF_ATTN_FWD, stream);
    kern<<<grid, THREADS, sl.total, stream>>>(p, sl, nitems, ngroups);
    SAT_CHECK_CUDA(cudaGetLastError());
    return SAT_OK;
}

template <typename T, bool PRECISE>
int launch_fwd(const FwdParams& p, cudaStream_t stream) {
#define SAT_FWD_CASE(DD, HH)                                                                           \
    if (p.D == DD && p.H == HH) {                                                                      \
        if (p.NB == 1) return launch_fwd_inst<T, DD, HH, 1, PRECISE>(p, stream);                       \
        if constexpr (DD == 512) return launch_fwd_inst<T, DD, HH, 5, PRECISE>(p, stream);             \
    }
    SAT_FWD_CASE(512, 512) SAT_FWD_CASE(512, 768) SAT_FWD_CASE(2048, 512) SAT_FWD_CASE(2048, 768)
#undef SAT_FWD_CASE
    sat_set_error("attention fwd (streaming): unsupported shape D=%d H=%d NB=%d", p.D, p.H, p.NB);
    return SAT_ERR_UNSUPPORTED;
}

// the streaming backward needs the per-step dpre buffer; the sequence driver asks this before carving it
template <typename T>
inline bool bwd_shape_eligible(int B, int L, int D, int H) { return shape_ok(L, D, H) && B >= MIN_ITEMS; }

template <typename T>
inline bool bwd_eligible(const BwdParams& p) {
    if (!p.dpre_out || p.da_acc || !bwd_shape_eligible<T>(p.B, p.L, p.D, p.H)) return false;
    const size_t oa = sizeof(T) == 2 ? 8 : 16;
    return aligned(p.a, 16) && aligned(p.wa, 16) && aligned(p.uf, 16) && (p.ld_uf % 4 == 0) && aligned(p.v, 16) &&
           aligned(p.dz, 16) && (p.ld_dz % 4 == 0) && (!p.dctx_extra || (aligned(p.dctx_extra, 16) && p.ld_dce % 4 == 0)) &&
           aligned(p.ctx, oa) && (p.ld_ctx % 4 == 0) && aligned(p.duf, oa) && (p.ld_duf % 4 == 0) && aligned(p.dpre_out, 16);
}

template <typename T, int D, int H, bool PRECISE>
int launch_bwd_inst(const BwdParams& p, cudaStream_t stream) {
    auto kern = bwd_stream_kernel<T, D, H, PRECISE>;
    static bool configured = false;
    if (!configured) {
        SAT_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_MAX));
        configured = true;
    }
    const BwdSmemS sl = bwd_smem<T, D, H>(p.L);
    SAT_REQUIRE(sl.nslots >= CWARPS && sl.total <= SMEM_MAX, "attention bwd (streaming): shared memory budget exceeded (L=%d)", p.L);
    const int grid = p.B < num_sms() ? p.B : num_sms();
    SatProfScope prof(SAT_PROF_ATTN_BWD, stream);
    kern<<<grid, THREADS, sl.total, stream>>>(p, sl);
    SAT_CHECK_CUDA(cudaGetLastError());
    return SAT_OK;
}

template <typename T, bool PRECISE>
int launch_bwd(const BwdParams& p, cudaStream_t stream) {
    if (p.D == 512 && p.H == 512) return launch_bwd_inst<T, 512, 512, PRECISE>(p, stream);
    if (p.D == 512 && p.H == 768) return launch_bwd_inst<T, 512, 768, PRECISE>(p, stream);
    if (p.D == 2048 && p.H == 512) return launch_bwd_inst<T, 2048, 512, PRECISE>(p, stream);
    if (p.D == 2048 && p.H == 768) return launch_bwd_inst<T, 2048, 768, PRECISE>(p, stream);
    sat_set_error("attention bwd (streaming): unsupported shape D=%d H=%d", p.D, p.H);
    return SAT_ERR_UNSUPPORTED;
}
}  // namespace attn_s

// dispatch: streaming fast path when the shape/alignment qualifies, generic cluster kernels otherwise
template <typename T, bool PRECISE>
int attn_step_fwd(const attn::FwdParams& p, cudaStream_t stream) {
    if (attn_s::fwd_eligible<T>(p)) return attn_s::launch_fwd<T, PRECISE>(p, stream);
    return attn::launch_fwd<T, PRECISE>(p, stream);
}
template <typename T, bool PRECISE>
int attn_step_bwd(const attn::BwdParams& p, cudaStream_t stream) {
    if (p.dpre_out) {
        SAT_REQUIRE(attn_s::bwd_eligible<T>(p), "attention bwd: dpre_out given but the streaming kernel does not support this shape/alignment");
        return attn_s::launch_bwd<T, PRECISE>(p, stream);
    }
    SAT_REQUIRE(p.dwa_acc != nullptr, "attention bwd: dwa_acc is NULL");
    return attn::launch_bwd<T, PRECISE>(p, stream);
}
