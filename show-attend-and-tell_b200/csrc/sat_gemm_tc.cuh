// bf16 tensor-core GEMM for sm_100a:  C[M,N] = A[M,K] * B[N,K]^T, fp32 accumulation in TMEM.
//   * operands staged global -> shared by TMA (cp.async.bulk.tensor, 128B swizzle) through a
//     multi-stage mbarrier ring,
//   * tcgen05.mma (cta_group::1, kind::f16, M=128, N=BN, K=16) issued by one elected thread,
//   * accumulator read back with tcgen05.ld (32x32b.x32) by four epilogue warps that apply a fused
//     epilogue functor (bias / activation / LSTM cell ...) and write straight to global memory.
// Either operand may be K-major (row = M/N index, K contiguous) or MN-major (row = K index, M/N
// contiguous); the latter serves dgrad/wgrad GEMMs without materialised transposes.
#pragma once
#include "sat_common.cuh"

namespace tc {
constexpr int BM = 128, BK = 64, UMMA_K = 16;
#ifndef SAT_GEMM_NPROD
#define SAT_GEMM_NPROD 2
#endif
// One thread can start a TMA copy only every ~60 ns and a wait -> expect_tx -> copy round costs ~250-300 ns (measured,
// tools/tma_bw.cu), which bounds a skinny GEMM's k-loop; so NPROD producer lanes take the k-blocks round-robin.
constexpr int NPROD = SAT_GEMM_NPROD;
}  // namespace tc
extern int g_gemm_wide;         // sat_set_gemm_wide_tile: 1 (default) = large K-major GEMMs use the 256 x 256 two-accumulator tile
extern int g_gemm_max_split;    // tuning knob (sat_set_gemm_splitk): largest split-K cluster size, 1 = never split
// Debug instrumentation (sat_gemm_phase_stamps): when the flag is set, CTA (0,0,0) of a GEMM whose epilogue carries a
// DBG_ID > 0 leaves %globaltimer stamps of its phases: 0 entry, 1 after the PDL wait, 2 first operand stage landed,
// 3 last MMA issued, 4 accumulator complete (epilogue released), 5 split-K exchange done, 6 epilogue done.
__device__ int g_gemm_dbg_on = 0;
__device__ unsigned long long g_gemm_dbg[4][8];
__device__ __forceinline__ unsigned long long gtimer() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
template <typename Epi> struct EpiDbgId { static constexpr int value = 0; };
// Epilogues that read / write several global arrays per element (LSTM cell, gate backward) opt in to the COALESCED drain: the
// accumulator tile is staged in the (dead) operand ring and re-read with consecutive lanes on consecutive 4-column chunks of
// one row, so a warp's global access covers whole 128-byte segments instead of 32 different rows (phase stamps of the in-loop
// GEMMs, profiles/r1_gemm_phases_*.txt: the row-per-thread drain was 7.8 of 15.4 us in the gate GEMM, 9.2 of 16.5 us in dz).
template <typename Epi> struct EpiCoalesce { static constexpr bool value = false; };
#define SAT_GEMM_STAMP(slot)                                                                         \
    do { if (EpiDbgId<Epi>::value > 0 && (blockIdx.x | blockIdx.y | blockIdx.z) == 0 && g_gemm_dbg_on)  \
             g_gemm_dbg[EpiDbgId<Epi>::value][slot] = gtimer(); } while (0)
namespace tc {
#ifndef SAT_GEMM_EPI_GROUPS
#define SAT_GEMM_EPI_GROUPS 4
#endif
// A warp keeps only a few global stores in flight (~2 GB/s per warp measured: with four epilogue warps the stores were
// 50-80% of the large GEMMs, profiles/r1_gemm_harness_epilogue_store_cost.txt), so the accumulator is drained by EPI_GROUPS
// groups of four warps (a warp may only touch the TMEM lane quarter warp % 4); the groups split the tile's columns.
constexpr int EPI_GROUPS = SAT_GEMM_EPI_GROUPS;
constexpr int EPI_WARPS = 4 * EPI_GROUPS;
constexpr int PROD_WARP0 = 2 + EPI_WARPS;      // first extra producer warp
// warp0 = TMA producer 0, warp1 = TMEM alloc + MMA issuer, warps 2 .. 2+EPI_WARPS-1 = epilogue, then further TMA producers
constexpr int THREADS = 32 * (2 + EPI_WARPS + (NPROD - 1));

template <int BN> struct Cfg {
    static constexpr int STAGES = (BN <= 128) ? 6 : 4;      // multiple of NPROD: a stage always has the same producer
    static constexpr int A_BYTES = BM * BK * 2;
    static constexpr int B_BYTES = BN * BK * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;
};

__device__ __forceinline__ uint64_t make_desc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr >> 4) & 0x3fff);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3fff) << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3fff) << 32;
    d |= (uint64_t)1 << 46;   // descriptor version 1 (Blackwell)
    d |= (uint64_t)2 << 61;   // SWIZZLE_128B
    return d;
}

__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float* v) {
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, float* v) {
    uint32_t r[8];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
        : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float* v) {
    uint32_t r[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
        "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}

// SPLIT > 1: split-K over a thread-block cluster of SPLIT CTAs along z (a skinny GEMM with few tiles but a long K would
// otherwise keep most SMs idle, each busy CTA pulling its operands at the ~130 GB/s one SM gets out of L2).  CTA z
// accumulates k-blocks [nk*z/SPLIT, nk*(z+1)/SPLIT) in TMEM; the partial tiles are exchanged through distributed shared
// memory reduce-scatter style -- CTA j receives columns [j*BN/SPLIT, (j+1)*BN/SPLIT) of every partial, adds them in a fixed
// order (deterministic) and runs the epilogue for those columns.
template <int BN, bool A_MN, bool B_MN, typename Epi, int SPLIT>
__global__ void __launch_bounds__(THREADS, 1)
gemm_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB,
            int M, int N, int K, Epi epi, int b_static) {
    using C = Cfg<BN>;
    extern __shared__ uint8_t smem_raw[];
    // 1024-byte alignment (128B swizzle atoms) as an OFFSET from the __shared__ array, so that the compiler still knows the
    // address space (a pointer rebuilt from an integer turns every access into a slow generic LD/ST)
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t* sA = smem;
    uint8_t* sB = smem + C::STAGES * C::A_BYTES;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + C::STAGES * C::STAGE_BYTES);
    uint64_t* empty_bar = full_bar + C::STAGES;
    uint64_t* tmem_full_bar = empty_bar + C::STAGES;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) SAT_GEMM_STAMP(0);
    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
    const int nk_all = (K + BK - 1) / BK;
    const int zsp = SPLIT > 1 ? (int)blockIdx.z : 0;
    const int kb0 = SPLIT > 1 ? nk_all * zsp / SPLIT : 0;
    const int nk = (SPLIT > 1 ? nk_all * (zsp + 1) / SPLIT : nk_all) - kb0;     // k-blocks of this CTA

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&mapA);
        tma_prefetch_desc(&mapB);
        for (int s = 0; s < C::STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        mbar_init(tmem_full_bar, 1);
        mbar_fence_init();
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     ::"r"(smem_u32(tmem_slot)), "n"(BN) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    if constexpr (SPLIT > 1) {                             // cluster barrier #1: the epilogue warps arrive once the accumulator is done
        if (warp < 2 || warp >= PROD_WARP0) cluster_arrive_release();
    }
    // An MN-major operand arrives as 64-column boxes (128-byte swizzle span), up to six copies per stage; one thread
    // can only start a copy every ~60 ns, so lane b of the producer warp issues box b of the stage.
    constexpr int A_BOXES = A_MN ? BM / 64 : 1, B_BOXES = B_MN ? BN / 64 : 1;
    static_assert(C::STAGES % NPROD == 0, "a stage must always be filled by the same producer");
    const bool producer = warp == 0 || warp >= PROD_WARP0;
    const int kb_first = warp == 0 ? 0 : warp - PROD_WARP0 + 1;
    auto load_b = [&](int kb, int s) {
        if (lane >= A_BOXES && lane < A_BOXES + B_BOXES) {
            const int c = lane - A_BOXES, k0 = (kb0 + kb) * BK;
            uint8_t* b_dst = sB + s * C::B_BYTES;
            if (b_static) {         // a weight matrix re-read every decode step: keep it in L2 across the attention kernel's stream
                if (!B_MN) tma_load_2d_hint(b_dst, &mapB, k0, n0, &full_bar[s], L2_EVICT_LAST);
                else tma_load_2d_hint(b_dst + c * (BK * 128), &mapB, n0 + c * 64, k0, &full_bar[s], L2_EVICT_LAST);
            } else {
                if (!B_MN) tma_load_2d(b_dst, &mapB, k0, n0, &full_bar[s]);
                else tma_load_2d(b_dst + c * (BK * 128), &mapB, n0 + c * 64, k0, &full_bar[s]);
            }
        }
    };
    // b_static: operand B (a packed weight matrix) is older than the previous kernel in the stream, so the first ring pass
    // of its tiles is requested BEFORE the programmatic-dependent-launch wait and lands while the predecessor drains.
    if (b_static && producer) {
        for (int kb = kb_first; kb < nk && kb < C::STAGES; kb += NPROD) {
            if (lane == 0) mbar_expect_tx(&full_bar[kb], C::STAGE_BYTES);
            __syncwarp();
            load_b(kb, kb);
        }
    }
    // Programmatic dependent launch: everything above (barriers, TMEM, descriptor prefetch) overlapped the previous
    // kernel's tail; its results are needed from here on.  Let the next kernel start its own prologue right away.
    pdl_wait();
    pdl_launch_dependents();
    if (threadIdx.x == 0) SAT_GEMM_STAMP(1);

    if (producer) {
        for (int kb = kb_first; kb < nk; kb += NPROD) {
            const int s = kb % C::STAGES;
            const uint32_t ph = (kb / C::STAGES) & 1;
            const bool b_done = b_static && kb < C::STAGES;       // expect_tx posted and B requested before the wait
            mbar_wait(&empty_bar[s], ph ^ 1);
            if (lane == 0 && !b_done) mbar_expect_tx(&full_bar[s], C::STAGE_BYTES);
            __syncwarp();
            uint8_t* a_dst = sA + s * C::A_BYTES;
            const int k0 = (kb0 + kb) * BK;
            if (lane < A_BOXES) {
                if (!A_MN) tma_load_2d(a_dst, &mapA, k0, m0, &full_bar[s]);
                else tma_load_2d(a_dst + lane * (BK * 128), &mapA, m0 + lane * 64, k0, &full_bar[s]);
            }
            if (!b_done) load_b(kb, s);
        }
    } else if (warp == 1) {
        if (elect_one()) {
            constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((A_MN ? 1u : 0u) << 15) |
                                       ((B_MN ? 1u : 0u) << 16) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
            for (int kb = 0; kb < nk; ++kb) {
                const int s = kb % C::STAGES;
                const uint32_t ph = (kb / C::STAGES) & 1;
                mbar_wait(&full_bar[s], ph);
                tc_fence_after();
                if (kb == 0) SAT_GEMM_STAMP(2);
                const uint32_t a_addr = smem_u32(sA + s * C::A_BYTES);
                const uint32_t b_addr = smem_u32(sB + s * C::B_BYTES);
#pragma unroll
                for (int kk = 0; kk < BK / UMMA_K; ++kk) {
                    // K-major: 16 elements = 32 B along the swizzled row; MN-major: 16 k-rows of 128 B
                    const uint64_t ad = A_MN ? make_desc(a_addr + kk * (UMMA_K * 128), BK * 128, 1024)
                                             : make_desc(a_addr + kk * (UMMA_K * 2), 16, 1024);
                    const uint64_t bd = B_MN ? make_desc(b_addr + kk * (UMMA_K * 128), BK * 128, 1024)
                                             : make_desc(b_addr + kk * (UMMA_K * 2), 16, 1024);
                    umma_bf16(tmem_base, ad, bd, idesc, (kb > 0 || kk > 0) ? 1u : 0u);
                }
                umma_commit(&empty_bar[s]);          // frees the smem stage when these MMAs retire
            }
            SAT_GEMM_STAMP(3);
            umma_commit(tmem_full_bar);              // accumulator complete
        }
    } else {
        const int q = warp & 3;                      // TMEM lane quarter this warp may access
        const int m = m0 + q * 32 + lane;
        const int grp = (warp - 2) >> 2;             // epilogue group of this warp
        {   // while the MMAs run: warm L2 with the epilogue's global inputs for the columns this CTA finishes
            constexpr int CWP = BN / SPLIT;
            constexpr int PWD = CWP / EPI_GROUPS >= 4 ? CWP / EPI_GROUPS : 4;
            const int pn = n0 + zsp * CWP + grp * PWD;
            if (grp * PWD < CWP && pn < N) epi.prefetch(m, pn, PWD);
        }
        mbar_wait(tmem_full_bar, 0);
        tc_fence_after();
        if (warp == 2 && lane == 0) SAT_GEMM_STAMP(4);
        if constexpr (SPLIT == 1 && EpiCoalesce<Epi>::value) {
            constexpr int XP = BN + 4;               // padded row pitch: float4 accesses of 8 consecutive rows hit distinct banks
            constexpr int PW0 = (BN / EPI_GROUPS > 8) ? BN / EPI_GROUPS : 8;
            constexpr int PW = PW0 > 32 ? 32 : PW0;  // staging piece: 8, 16 or 32 columns
            static_assert(BN % PW == 0 && BM * XP * 4 <= C::STAGES * C::STAGE_BYTES, "staging buffer must fit in the stage ring");
            float* xbuf = reinterpret_cast<float*>(smem);     // the accumulator is complete: every MMA has read its stage
            const int row = q * 32 + lane;
#pragma unroll 1
            for (int p = grp; p < BN / PW; p += EPI_GROUPS) {
                float v[PW];
                const uint32_t ta = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(p * PW);
                if constexpr (PW == 32) tmem_ld32(ta, v); else if constexpr (PW == 16) tmem_ld16(ta, v); else tmem_ld8(ta, v);
                float4* dst = reinterpret_cast<float4*>(xbuf + (size_t)row * XP + p * PW);
#pragma unroll
                for (int c4 = 0; c4 < PW / 4; ++c4) dst[c4] = make_float4(v[4 * c4], v[4 * c4 + 1], v[4 * c4 + 2], v[4 * c4 + 3]);
            }
            asm volatile("bar.sync 2, %0;" ::"n"(EPI_WARPS * 32) : "memory");      // epilogue warps only
            constexpr int C4 = BN / 4;
            for (int c = (warp - 2) * 32 + lane; c < BM * C4; c += EPI_WARPS * 32) {
                const int r = c / C4, c4 = c - r * C4;
                const float4 t = *reinterpret_cast<const float4*>(xbuf + (size_t)r * XP + c4 * 4);
                const float acc[4] = {t.x, t.y, t.z, t.w};
                const int n = n0 + c4 * 4;
                if (n < N) epi.template apply<4>(m0 + r, n, acc);
            }
        } else if constexpr (SPLIT == 1) {
            constexpr int NC = BN / 32;
            if constexpr (NC >= EPI_GROUPS) {        // whole 32-column chunks, dealt round-robin to the groups
#pragma unroll 1
                for (int c = grp; c < NC; c += EPI_GROUPS) {
                    float v[32];
                    tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(c * 32), v);
                    if (n0 + c * 32 < N) epi.template apply<32>(m, n0 + c * 32, v);
                }
            } else {                                 // narrow tile: each group takes BN / EPI_GROUPS columns
                constexpr int W = BN / EPI_GROUPS;
                static_assert(W == 8 || W == 16, "narrow tiles are split into 8- or 16-column pieces");
                float v[W];
                const uint32_t ta = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(grp * W);
                if constexpr (W == 16) tmem_ld16(ta, v); else tmem_ld8(ta, v);
                if (n0 + grp * W < N) epi.template apply<W>(m, n0 + grp * W, v);
            }
        } else {
            // Exchange buffer [SPLIT][BM][XP] fp32 ALIASES the operand stages: a CTA signals barrier #1 only after its own
            // accumulator is complete (all its MMAs have read their stages), so once #1 completes every peer's stages are dead.
            constexpr int CW = BN / SPLIT;           // columns this CTA finishes
            constexpr int XP = CW + 4;               // padded row pitch: float4 accesses of 8 consecutive rows hit distinct banks
            constexpr int PW0 = (BN / EPI_GROUPS > 8) ? BN / EPI_GROUPS : 8;
            constexpr int PW = (PW0 > CW ? CW : PW0) > 32 ? 32 : (PW0 > CW ? CW : PW0);      // scatter piece: 8, 16 or 32 columns
            constexpr int RW0 = (CW / EPI_GROUPS > 4) ? CW / EPI_GROUPS : 4;
            constexpr int RW = RW0 > 32 ? 32 : RW0;                                            // reduce piece: 4 .. 32 columns
            static_assert(PW == 8 || PW == 16 || PW == 32, "scatter pieces are 8, 16 or 32 columns");
            static_assert(CW % PW == 0 && CW % RW == 0 && RW % 4 == 0, "column slices must divide evenly");
            static_assert(SPLIT * BM * XP * 4 <= C::STAGES * C::STAGE_BYTES, "exchange buffer must fit in the stage ring");
            float* xbuf = reinterpret_cast<float*>(smem);
            const int row = q * 32 + lane;
            __syncwarp();
            cluster_arrive_release();                // #1: my accumulator is complete, my stages may be overwritten
            cluster_wait_acquire();
#pragma unroll 1
            for (int p = grp; p < BN / PW; p += EPI_GROUPS) {
                float v[PW];
                const uint32_t ta = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(p * PW);
                if constexpr (PW == 32) tmem_ld32(ta, v); else if constexpr (PW == 16) tmem_ld16(ta, v); else tmem_ld8(ta, v);
                const int col = p * PW;
                const uint32_t dst = dsmem_addr(xbuf + ((size_t)zsp * BM + row) * XP + (col % CW), (uint32_t)(col / CW));
#pragma unroll
                for (int c4 = 0; c4 < PW / 4; ++c4)
                    dsmem_st_f32x4(dst + 16 * c4, v[4 * c4], v[4 * c4 + 1], v[4 * c4 + 2], v[4 * c4 + 3]);
            }
            __syncwarp();
            cluster_arrive_release();                // #2: my slices are in every peer's buffer ...
            cluster_wait_acquire();                  //     ... and every peer's slices are in mine
            if (warp == 2 && lane == 0) SAT_GEMM_STAMP(5);
            if constexpr (EpiCoalesce<Epi>::value) {
                constexpr int C4 = CW / 4;
                for (int c = (warp - 2) * 32 + lane; c < BM * C4; c += EPI_WARPS * 32) {
                    const int r = c / C4, c4 = c - r * C4;
                    float acc[4] = {0.0f, 0.0f, 0.0f, 0.0f};
#pragma unroll
                    for (int sp = 0; sp < SPLIT; ++sp) {             // fixed order: deterministic
                        const float4 t = *reinterpret_cast<const float4*>(xbuf + ((size_t)sp * BM + r) * XP + c4 * 4);
                        acc[0] += t.x; acc[1] += t.y; acc[2] += t.z; acc[3] += t.w;
                    }
                    const int n = n0 + zsp * CW + c4 * 4;
                    if (n < N) epi.template apply<4>(m0 + r, n, acc);
                }
            } else
#pragma unroll 1
            for (int r = grp; r < CW / RW; r += EPI_GROUPS) {
                float acc[RW];
#pragma unroll
                for (int i = 0; i < RW; ++i) acc[i] = 0.0f;
#pragma unroll
                for (int sp = 0; sp < SPLIT; ++sp) {                 // fixed order: deterministic
                    const float4* src = reinterpret_cast<const float4*>(xbuf + ((size_t)sp * BM + row) * XP + r * RW);
#pragma unroll
                    for (int c4 = 0; c4 < RW / 4; ++c4) {
                        const float4 t = src[c4];
                        acc[4 * c4] += t.x; acc[4 * c4 + 1] += t.y; acc[4 * c4 + 2] += t.z; acc[4 * c4 + 3] += t.w;
                    }
                }
                const int n = n0 + zsp * CW + r * RW;
                if (n < N) epi.template apply<RW>(m, n, acc);
            }
        }
    }
    if constexpr (SPLIT > 1) {
        if (warp < 2 || warp >= PROD_WARP0) {        // the other warps (arrived at #1 after setup) take part in both barriers too
            __syncwarp();
            cluster_wait_acquire();
            cluster_arrive_release();
            cluster_wait_acquire();
        }
    }
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x == 0) SAT_GEMM_STAMP(6);
    if (warp == 1) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(BN) : "memory");
    }
}

// ---- wide tile: 256 x 256 per CTA (two 128-row accumulators sharing one B tile) --------------------------------------------
// The 128 x 256 tile above pulls 48 KB from L2 per k-block for 0.26 us of MMAs = 185 GB/s per SM, 27 TB/s over 148 SMs -- more
// than L2 delivers (~12 TB/s): ncu shows K1 (M = B L = 50 176, N = 512, K = 2048) at 47.7 % tensor-pipe activity = the L2 bound of
// that tile shape (85 flop per operand byte).  Here one CTA keeps TWO M-tiles' accumulators in TMEM (2 x 256 = all 512 columns) and
// issues both tiles' MMAs against the SAME B stage: 64 KB per k-block for twice the math (131 flop per byte, the intensity a
// cta_group::2 pair would have) with nothing shared between SMs.  K-major operands only, BN = 256, no split-K; three 64 KB stages;
// one producer thread (a k-block lasts 0.52 us, a TMA round ~0.3 us); the sixteen epilogue warps drain the two accumulators one
// after the other through the coalesced staging path.
struct Cfg2M {
    static constexpr int BN = 256, MT = 2, STAGES = 3;
    static constexpr int A_BYTES = MT * BM * BK * 2;     // 32 KB: 256 rows x 128 B, m-tile j starts at j * 16 KB
    static constexpr int B_BYTES = BN * BK * 2;          // 32 KB
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 + 256;
};

template <typename Epi>
__global__ void __launch_bounds__(THREADS, 1)
gemm2m_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB, int M, int N, int K, Epi epi) {
    using C = Cfg2M;
    constexpr int BN = C::BN;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t* sA = smem;
    uint8_t* sB = smem + C::STAGES * C::A_BYTES;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + C::STAGES * C::STAGE_BYTES);
    uint64_t* empty_bar = full_bar + C::STAGES;
    uint64_t* tmem_full_bar = empty_bar + C::STAGES;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m0 = blockIdx.y * (C::MT * BM), n0 = blockIdx.x * BN;
    const int nk = (K + BK - 1) / BK;
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&mapA);
        tma_prefetch_desc(&mapB);
        for (int s = 0; s < C::STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        mbar_init(tmem_full_bar, 1);
        mbar_fence_init();
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "n"(C::MT * BN) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();
    pdl_launch_dependents();
    if (warp == 0) {
        if (lane == 0) {
            for (int kb = 0; kb < nk; ++kb) {
                const int s = kb % C::STAGES;
                mbar_wait(&empty_bar[s], ((kb / C::STAGES) & 1) ^ 1);
                mbar_expect_tx(&full_bar[s], C::STAGE_BYTES);
                tma_load_2d(sA + s * C::A_BYTES, &mapA, kb * BK, m0, &full_bar[s]);      // 256 rows: both M-tiles
                tma_load_2d(sB + s * C::B_BYTES, &mapB, kb * BK, n0, &full_bar[s]);
            }
        }
    } else if (warp == 1) {
        if (elect_one()) {
            constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
            for (int kb = 0; kb < nk; ++kb) {
                const int s = kb % C::STAGES;
                mbar_wait(&full_bar[s], (kb / C::STAGES) & 1);
                tc_fence_after();
                const uint32_t a_addr = smem_u32(sA + s * C::A_BYTES), b_addr = smem_u32(sB + s * C::B_BYTES);
#pragma unroll
                for (int kk = 0; kk < BK / UMMA_K; ++kk) {
                    const uint64_t bd = make_desc(b_addr + kk * (UMMA_K * 2), 16, 1024);
#pragma unroll
                    for (int j = 0; j < C::MT; ++j)
                        umma_bf16(tmem_base + (uint32_t)(j * BN), make_desc(a_addr + j * (BM * BK * 2) + kk * (UMMA_K * 2), 16, 1024), bd, idesc,
                                  (kb > 0 || kk > 0) ? 1u : 0u);
                }
                umma_commit(&empty_bar[s]);
            }
            umma_commit(tmem_full_bar);
        }
    } else if (warp < 2 + EPI_WARPS) {
        const int q = warp & 3, grp = (warp - 2) >> 2;
        mbar_wait(tmem_full_bar, 0);
        tc_fence_after();
        constexpr int XP = BN + 4, PW = 32, C4 = BN / 4;
        static_assert(BM * XP * 4 <= C::STAGES * C::STAGE_BYTES, "staging buffer must fit in the stage ring");
        float* xbuf = reinterpret_cast<float*>(smem);
        const int row = q * 32 + lane;
#pragma unroll 1
        for (int j = 0; j < C::MT; ++j) {
            if (j > 0) asm volatile("bar.sync 2, %0;" ::"n"(EPI_WARPS * 32) : "memory");      // everyone has read the previous tile
#pragma unroll 1
            for (int p = grp; p < BN / PW; p += EPI_GROUPS) {
                float v[PW];
                tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(j * BN + p * PW), v);
                float4* dst = reinterpret_cast<float4*>(xbuf + (size_t)row * XP + p * PW);
#pragma unroll
                for (int c4 = 0; c4 < PW / 4; ++c4) dst[c4] = make_float4(v[4 * c4], v[4 * c4 + 1], v[4 * c4 + 2], v[4 * c4 + 3]);
            }
            asm volatile("bar.sync 2, %0;" ::"n"(EPI_WARPS * 32) : "memory");
            for (int c = (warp - 2) * 32 + lane; c < BM * C4; c += EPI_WARPS * 32) {
                const int r = c / C4, c4 = c - r * C4;
                const float4 t = *reinterpret_cast<const float4*>(xbuf + (size_t)r * XP + c4 * 4);
                const float acc[4] = {t.x, t.y, t.z, t.w};
                const int n = n0 + c4 * 4;
                if (n < N) epi.template apply<4>(m0 + j * BM + r, n, acc);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(C::MT * BN) : "memory");
    }
}

// ---- host side ---------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
inline EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

inline int make_map_uncached(CUtensorMap* map, const void* ptr, uint64_t inner, uint64_t outer, int64_t ld,
                             uint32_t box_inner, uint32_t box_outer);
// Tensor maps are a pure function of (pointer, shape, pitch, box, element type): a small per-thread direct-mapped cache saves the
// driver call for the ~90 GEMMs of an eagerly launched train step, whose operands live at fixed workspace addresses (under CUDA-graph
// replay the maps are baked into the captured kernel parameters and nothing is encoded at all).
struct MapKey { const void* ptr; uint64_t inner, outer; int64_t ld; uint32_t bi, bo, esz; };
struct MapSlot { MapKey k; CUtensorMap m; bool ok; };
inline MapSlot* map_slot(const MapKey& k) {
    static thread_local MapSlot cache[256] = {};
    uint64_t h = ((uint64_t)(uintptr_t)k.ptr >> 7) * 0x9E3779B97F4A7C15ull ^ (k.inner * 131 + k.outer) * 0xC2B2AE3D27D4EB4Full ^ ((uint64_t)k.bo << 20 | k.bi << 4 | k.esz);
    return &cache[(h >> 32) & 255];
}
inline bool map_key_eq(const MapKey& a, const MapKey& b) {
    return a.ptr == b.ptr && a.inner == b.inner && a.outer == b.outer && a.ld == b.ld && a.bi == b.bi && a.bo == b.bo && a.esz == b.esz;
}

// 2-D bf16 tensor map: `inner` contiguous elements per row, `outer` rows, row pitch ld elements.
inline int make_map(CUtensorMap* map, const void* ptr, uint64_t inner, uint64_t outer, int64_t ld,
                    uint32_t box_inner, uint32_t box_outer) {
    const MapKey key{ptr, inner, outer, ld, box_inner, box_outer, 2u};
    MapSlot* slot = map_slot(key);
    if (slot->ok && map_key_eq(slot->k, key)) { *map = slot->m; return SAT_OK; }
    SAT_TRY(make_map_uncached(map, ptr, inner, outer, ld, box_inner, box_outer));
    slot->k = key; slot->m = *map; slot->ok = true;
    return SAT_OK;
}
inline int make_map_uncached(CUtensorMap* map, const void* ptr, uint64_t inner, uint64_t outer, int64_t ld,
                             uint32_t box_inner, uint32_t box_outer) {
    EncodeTiledFn fn = get_encode_fn();
    if (!fn) { sat_set_error("cuTensorMapEncodeTiled entry point not found"); return SAT_ERR_CUDA; }
    if (((uintptr_t)ptr & 15) || (ld % 8)) {
        sat_set_error("tcgen05 GEMM operand needs 16B-aligned base and ld %% 8 == 0 (ptr=%p ld=%lld)", ptr, (long long)ld);
        return SAT_ERR_ARG;
    }
    cuuint64_t dims[2] = {inner, outer};
    cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
    cuuint32_t box[2] = {box_inner, box_outer};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { sat_set_error("cuTensorMapEncodeTiled failed: %d", (int)r); return SAT_ERR_CUDA; }
    return SAT_OK;
}

template <int BN, bool A_MN, bool B_MN, typename Epi, int SPLIT>
int launch_bn(cudaStream_t stream, const CUtensorMap& mA, const CUtensorMap& mB, int M, int N, int K, const Epi& epi, int b_static) {
    auto kern = gemm_kernel<BN, A_MN, B_MN, Epi, SPLIT>;
    constexpr int smem = Cfg<BN>::SMEM_BYTES;
    static SatPerDevice configured;
    SAT_TRY(configured.once([&]() -> int {
        SAT_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        return SAT_OK;
    }));
    dim3 grid(sat_cdiv(N, BN), sat_cdiv(M, BM), SPLIT);
    cudaLaunchConfig_t cfg;
    cudaLaunchAttribute attr[2];
    pdl_config(&cfg, attr, grid, dim3(THREADS), smem, stream);
    if (SPLIT > 1) {
        attr[1].id = cudaLaunchAttributeClusterDimension;
        attr[1].val.clusterDim.x = 1; attr[1].val.clusterDim.y = 1; attr[1].val.clusterDim.z = SPLIT;
        cfg.numAttrs = 2;
    }
    SAT_CHECK_CUDA(cudaLaunchKernelEx(&cfg, kern, mA, mB, M, N, K, epi, b_static));
    return SAT_OK;
}

// Tile width and split-K factor from a cost model fitted to tools/gemm_bench.cu on B200 (profiles/r1_gemm_harness_*.txt), in
// units of ~6.5 ns:  waves x (launch 560 + exchange [150 + BN x split] + drain 2.4 x BN / split + k-blocks per CTA x kb(BN)),
// kb = 20 / 24 / 26 / 48 for BN = 32 / 64 / 128 / 256 (0.155 us is the floor of a k-block up to BN = 128).  It keeps the per-step
// skinny GEMMs (M = batch) in ONE wave of <= #SMs CTAs, splits K over a cluster when the tiles alone would leave most SMs idle
// (dh: 16 tiles x 40 k-blocks; dW_att: 64 tiles x 784 k-blocks) and takes the widest tile that does not add a wave otherwise.
inline void pick_tile(int M, int N, int K, bool b_mn, int* bn_out, int* split_out, int sm_budget = 0) {
    static int dev_sms = 0;
    if (!dev_sms) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&dev_sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) dev_sms = 148;
    }
    // sm_budget > 0: the SMs this GEMM can count on while another kernel holds the rest (half-batch pipelining)
    const int sms = (sm_budget > 0 && sm_budget < dev_sms) ? sm_budget : dev_sms;
    const int mt = sat_cdiv(M, BM), nk = sat_cdiv(K, BK);
    const int cands[4] = {256, 128, 64, 32};
    int best = b_mn ? 64 : 32, best_split = 1;
    double best_cost = 1e30;
    for (int i = 0; i < 4; ++i) {
        const int bn = cands[i];
        if (b_mn && bn < 64) continue;
        if (bn > 32 && bn / 2 >= N && !(b_mn && bn == 64)) continue;   // tile far wider than N
        const long long tiles = (long long)mt * sat_cdiv(N, bn);
        const double stage_kb = bn == 256 ? 48.0 : bn == 128 ? 26.0 : bn == 64 ? 24.0 : 20.0;
        for (int split = 1; split <= 4 && split <= g_gemm_max_split; split *= 2) {
            if (split > 1 && (bn < 64 || tiles * split > sms || nk / split < 4)) continue;
            const double waves = (double)((tiles * split + sms - 1) / sms);
            const double cost = waves * (560.0 + (split > 1 ? 150.0 + (double)bn * split : 0.0) + 2.4 * bn / split +
                                         sat_cdiv(nk, split) * stage_kb);
            if (cost < best_cost * 0.999) { best_cost = cost; best = bn; best_split = split; }
        }
    }
    *bn_out = best; *split_out = best_split;
}

template <bool A_MN, bool B_MN, typename Epi>
int launch_mn(cudaStream_t stream, const bf16* A, int64_t lda, const bf16* B, int64_t ldb, int M, int N, int K,
              const Epi& epi, int bn, int split, int b_static) {
    CUtensorMap mA, mB;
    if (!A_MN) SAT_TRY(make_map(&mA, A, K, M, lda, 64, BM));
    else       SAT_TRY(make_map(&mA, A, M, K, lda, 64, BK));
    const int box_n = bn;
    if (!B_MN) SAT_TRY(make_map(&mB, B, K, N, ldb, 64, box_n));
    else       SAT_TRY(make_map(&mB, B, N, K, ldb, 64, BK));
#define SAT_TC_SPLITS(BN_)                                                                                     \
        case BN_:                                                                                              \
            if (split == 4) return launch_bn<BN_, A_MN, B_MN, Epi, 4>(stream, mA, mB, M, N, K, epi, b_static);           \
            if (split == 2) return launch_bn<BN_, A_MN, B_MN, Epi, 2>(stream, mA, mB, M, N, K, epi, b_static);           \
            return launch_bn<BN_, A_MN, B_MN, Epi, 1>(stream, mA, mB, M, N, K, epi, b_static);
    switch (bn) {
        SAT_TC_SPLITS(256)
        SAT_TC_SPLITS(128)
        SAT_TC_SPLITS(64)
        case 32:
            if constexpr (!B_MN) return launch_bn<32, A_MN, B_MN, Epi, 1>(stream, mA, mB, M, N, K, epi, b_static);
        default: break;
    }
#undef SAT_TC_SPLITS
    sat_set_error("bad BN %d", bn);
    return SAT_ERR_ARG;
}

template <typename Epi>
int launch_2m(cudaStream_t stream, const bf16* A, int64_t lda, const bf16* B, int64_t ldb, int M, int N, int K, const Epi& epi) {
    CUtensorMap mA, mB;
    SAT_TRY(make_map(&mA, A, K, M, lda, 64, Cfg2M::MT * BM));
    SAT_TRY(make_map(&mB, B, K, N, ldb, 64, Cfg2M::BN));
    auto kern = gemm2m_kernel<Epi>;
    static SatPerDevice configured;
    SAT_TRY(configured.once([&]() -> int {
        SAT_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg2M::SMEM_BYTES));
        return SAT_OK;
    }));
    cudaLaunchConfig_t cfg;
    cudaLaunchAttribute attr[1];
    pdl_config(&cfg, attr, dim3(sat_cdiv(N, Cfg2M::BN), sat_cdiv(M, Cfg2M::MT * BM)), dim3(THREADS), Cfg2M::SMEM_BYTES, stream);
    SAT_CHECK_CUDA(cudaLaunchKernelEx(&cfg, kern, mA, mB, M, N, K, epi));
    return SAT_OK;
}

template <typename Epi>
int launch(cudaStream_t stream, const bf16* A, int64_t lda, bool a_kmajor, const bf16* B, int64_t ldb, bool b_kmajor,
           int M, int N, int K, const Epi& epi, int bn_override = 0, int split_override = 0, int sm_budget = 0, int b_static = 0) {
    if (M <= 0 || N <= 0) return SAT_OK;
    SAT_REQUIRE(K > 0, "tcgen05 GEMM needs K > 0");
    if constexpr (EpiCoalesce<Epi>::value) {
        // wide tile where the grid still fills the machine at least twice over (K1 at c2: 196 x 2 CTAs); bn_override = 512 forces it
        if (a_kmajor && b_kmajor && !sm_budget && (bn_override == 512 || (g_gemm_wide && !bn_override && N >= 256 && K >= 256 &&
            (long long)sat_cdiv(M, Cfg2M::MT * BM) * sat_cdiv(N, Cfg2M::BN) >= 2 * 148)))
            return launch_2m<Epi>(stream, A, lda, B, ldb, M, N, K, epi);
    }
    SAT_REQUIRE(bn_override != 512, "the 256 x 256 tile needs K-major operands and a coalescing epilogue");
    int bn = 0, split = 1;
    pick_tile(M, N, K, !b_kmajor, &bn, &split, sm_budget);
    if (bn_override) { bn = bn_override; split = split_override > 0 ? split_override : 1; }
    if (a_kmajor && b_kmajor) return launch_mn<false, false, Epi>(stream, A, lda, B, ldb, M, N, K, epi, bn, split, b_static);
    if (a_kmajor && !b_kmajor) return launch_mn<false, true, Epi>(stream, A, lda, B, ldb, M, N, K, epi, bn, split, b_static);
    if (!a_kmajor && b_kmajor) return launch_mn<true, false, Epi>(stream, A, lda, B, ldb, M, N, K, epi, bn, split, b_static);
    return launch_mn<true, true, Epi>(stream, A, lda, B, ldb, M, N, K, epi, bn, split, b_static);
}
}  // namespace tc
