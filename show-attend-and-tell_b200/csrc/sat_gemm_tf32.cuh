// fp32-accurate GEMM on the 5th-gen tensor cores ("3xTF32"):  C[M,N] = A[M,K] * B[N,K]^T, both operands fp32 and K-major.
//
// The fp32 compute mode (config c1, bit-exact beam search c4) may not round its GEMM inputs to bf16, and a SIMT FMA kernel
// tops out near 25 TFLOP/s on this part (r2 beam timeline: the three fp32 GEMMs were 14 of 22 ms of a 128-image beam-5
// search).  tcgen05.mma.kind::tf32 reads only the top 19 bits of an fp32 operand, so every operand element is split
//     x = hi + lo,   hi = tf32(x) (cvt.rna: round to nearest, low 13 bits zero),   lo = x - hi  (exact in fp32)
// and a product is rebuilt from three MMAs into one TMEM accumulator:  lo_a*hi_b + hi_a*lo_b + hi_a*hi_b.  The dropped
// lo_a*lo_b term and the truncation of lo inside the MMA are each <= 2^-21 relative to |a b| and zero-mean: the result is
// within a few fp32 ulps of an exact-fp32 dot product, far inside what re-ordering an fp32 sum already moves.
//
// Structure (one CTA = one 128 x BN tile, BK = 32 fp32 = one 128-byte swizzle row per k-block):
//   warp 0        TMA producer: raw fp32 tiles of A and B into the stage's "hi" buffers (cp.async.bulk.tensor, 128B swizzle)
//   warps 2..17   SPLITTERS while the k-loop runs: rewrite each landed tile in place as hi and write lo to the twin buffer
//                 (an elementwise pass: the swizzled layout is preserved without ever being decoded), then
//                 fence.proxy.async + arrive on the stage's split barrier;  afterwards the same warps are the EPILOGUE:
//                 TMEM -> registers -> shared (coalescing stage) -> fused epilogue functor -> global
//   warp 1        allocates TMEM and, one elected thread, issues the 12 MMAs (4 k-steps x 3 terms) of every k-block
// Out-of-range rows / k are zero-filled by TMA and contribute nothing.
#pragma once
#include "sat_common.cuh"
#include "sat_gemm_tc.cuh"

extern int g_fp32_gemm_mode;     // sat_set_fp32_gemm_mode: 1 (default) = 3xTF32 tensor-core kernel where it applies, 0 = SIMT always

namespace tc32 {
constexpr int BM = 128, BK = 32, UMMA_K = 8;
constexpr int WORK_WARPS = 16;                         // splitters, then epilogue (4 groups x 4 TMEM lane quarters)
constexpr int THREADS = 32 * (2 + WORK_WARPS);
constexpr int A_BYTES = BM * BK * 4;                    // 16 KB
// The tensor core adds every MMA's products into the fp32 accumulator with TRUNCATION, so the error of one long accumulation chain
// grows linearly with its length (measured with one accumulator: 6e-4 absolute on sums of magnitude ~12 at K = 2560, i.e. 5e-5
// relative -- 100x what 3xTF32's input split leaves).  Hence NACC accumulators in TMEM (4 x BN <= 512 columns): the hi*hi
// products of the four k-steps of a k-block are dealt round-robin to three of them (each chain a third as long) and the two
// small cross terms go to the fourth (their truncation is relative to a 2^-11 times smaller sum); the epilogue adds the four
// in fp32 with round-to-nearest.
constexpr int NACC = 4;

template <int BN> struct Cfg {
    static constexpr int B_BYTES = BN * BK * 4;
    static constexpr int STAGE_BYTES = 2 * (A_BYTES + B_BYTES);     // [A hi | A lo | B hi | B lo]
    static constexpr int STAGES = (200 * 1024) / STAGE_BYTES;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;
    static_assert(STAGES >= 3, "need at least three stages");
    static_assert(BN == 32 || BN == 64 || BN == 128, "NACC * BN TMEM columns must be a power of two <= 512");
};

__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ float tf32_rna(float x) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return __uint_as_float(r);
}

// BSPLIT: the B operand (weights that serve many launches, e.g. the 51 decode steps of a beam search) arrives already split --
// mapB = tf32(B), mapBl = B - tf32(B), made once by split_operand() -- so the work warps split only the A tile.  A k-block of the
// unsplit kernel moves ~192 KB through shared memory (split: 32 KB read + 64 KB written; MMAs: 3 x (A + B) read), which is what
// bounds it at ~1 us; the pre-split B takes a quarter of that away.
template <int BN, typename Epi, bool BSPLIT>
__global__ void __launch_bounds__(THREADS, 1)
gemm_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB, const __grid_constant__ CUtensorMap mapBl,
            int M, int N, int K, Epi epi) {
    using C = Cfg<BN>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + C::STAGES * C::STAGE_BYTES);
    uint64_t* split_bar = full_bar + C::STAGES;
    uint64_t* empty_bar = split_bar + C::STAGES;
    uint64_t* tmem_full_bar = empty_bar + C::STAGES;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
    const int nk = (K + BK - 1) / BK;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&mapA);
        tma_prefetch_desc(&mapB);
        if constexpr (BSPLIT) tma_prefetch_desc(&mapBl);
        for (int s = 0; s < C::STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&split_bar[s], WORK_WARPS); mbar_init(&empty_bar[s], 1); }
        mbar_init(tmem_full_bar, 1);
        mbar_fence_init();
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "n"(NACC * BN) : "memory");
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
                const uint32_t ph = (kb / C::STAGES) & 1;
                mbar_wait(&empty_bar[s], ph ^ 1);
                uint8_t* st = smem + s * C::STAGE_BYTES;
                mbar_expect_tx(&full_bar[s], A_BYTES + (BSPLIT ? 2 : 1) * C::B_BYTES);
                tma_load_2d(st, &mapA, kb * BK, m0, &full_bar[s]);                            // raw A -> "A hi"
                tma_load_2d(st + 2 * A_BYTES, &mapB, kb * BK, n0, &full_bar[s]);              // raw B (or its hi part) -> "B hi"
                if constexpr (BSPLIT) tma_load_2d(st + 2 * A_BYTES + C::B_BYTES, &mapBl, kb * BK, n0, &full_bar[s]);      // -> "B lo"
            }
        }
    } else if (warp == 1) {
        if (elect_one()) {
            // instruction descriptor: D = fp32 (bit 4), A = B = TF32 (format 2 at bits 7 and 10), both K-major, N >> 3, M >> 4
            constexpr uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
            for (int kb = 0; kb < nk; ++kb) {
                const int s = kb % C::STAGES;
                const uint32_t ph = (kb / C::STAGES) & 1;
                mbar_wait(&split_bar[s], ph);
                tc_fence_after();
                const uint32_t a_hi = smem_u32(smem + s * C::STAGE_BYTES), a_lo = a_hi + A_BYTES;
                const uint32_t b_hi = a_hi + 2 * A_BYTES, b_lo = b_hi + C::B_BYTES;
#pragma unroll
                for (int kk = 0; kk < BK / UMMA_K; ++kk) {
                    const uint32_t o = kk * (UMMA_K * 4);                                     // 8 fp32 = 32 B along the swizzled row
                    const uint64_t ah = tc::make_desc(a_hi + o, 16, 1024), al = tc::make_desc(a_lo + o, 16, 1024);
                    const uint64_t bh = tc::make_desc(b_hi + o, 16, 1024), bl = tc::make_desc(b_lo + o, 16, 1024);
                    // hi*hi of k-step number `step` goes to accumulator step % 3, the two small terms to accumulator 3 (see NACC)
                    const int step = kb * (BK / UMMA_K) + kk, j = step % (NACC - 1);
                    umma_tf32(tmem_base + (uint32_t)((NACC - 1) * BN), al, bh, idesc, step > 0 ? 1u : 0u);
                    umma_tf32(tmem_base + (uint32_t)((NACC - 1) * BN), ah, bl, idesc, 1u);
                    umma_tf32(tmem_base + (uint32_t)(j * BN), ah, bh, idesc, step >= NACC - 1 ? 1u : 0u);
                }
                tc::umma_commit(&empty_bar[s]);
            }
            tc::umma_commit(tmem_full_bar);
        }
    } else {
        const int wt = threadIdx.x - 64;                 // 0 .. 511
        // ---- split the landed tiles: hi in place, lo to the twin buffer ----
        constexpr int A_CH = A_BYTES / 16, B_CH = C::B_BYTES / 16, SPLIT_CH = BSPLIT ? A_CH : A_CH + B_CH;
        for (int kb = 0; kb < nk; ++kb) {
            const int s = kb % C::STAGES;
            const uint32_t ph = (kb / C::STAGES) & 1;
            mbar_wait(&full_bar[s], ph);
            float4* ahi = reinterpret_cast<float4*>(smem + s * C::STAGE_BYTES);
            float4* alo = ahi + A_CH;
            float4* bhi = alo + A_CH;
            float4* blo = bhi + B_CH;
#pragma unroll 2
            for (int c = wt; c < SPLIT_CH; c += WORK_WARPS * 32) {
                float4* hp = c < A_CH ? ahi + c : bhi + (c - A_CH);
                float4* lp = c < A_CH ? alo + c : blo + (c - A_CH);
                const float4 x = *hp;
                float4 h, l;
                h.x = tf32_rna(x.x); h.y = tf32_rna(x.y); h.z = tf32_rna(x.z); h.w = tf32_rna(x.w);
                l.x = x.x - h.x; l.y = x.y - h.y; l.z = x.z - h.z; l.w = x.w - h.w;
                *hp = h; *lp = l;
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");      // generic-proxy writes -> visible to the MMA's async proxy
            __syncwarp();
            if (lane == 0) mbar_arrive(&split_bar[s]);
        }
        // ---- epilogue (coalesced drain through the dead stage ring, as in sat_gemm_tc.cuh) ----
        // a warp may only touch the TMEM lane quarter (warp index in the CTA) % 4; warps 2..17 = four groups covering all quarters
        const int q = warp & 3, grp = (warp - 2) >> 2;
        mbar_wait(tmem_full_bar, 0);
        tc_fence_after();
        constexpr int XP = BN + 4;
        constexpr int PW = BN / 4 > 32 ? 32 : BN / 4;     // 4 groups split the columns: 8, 16 or 32 per piece
        static_assert(BM * XP * 4 <= C::STAGES * C::STAGE_BYTES, "staging buffer must fit in the stage ring");
        float* xbuf = reinterpret_cast<float*>(smem);
        const int row = q * 32 + lane;
#pragma unroll 1
        for (int p = grp; p < BN / PW; p += 4) {
            float v[PW], u[PW];
            const uint32_t ta = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(p * PW);
            auto ld = [&](int acc, float* o) {
                const uint32_t t = ta + (uint32_t)(acc * BN);
                if constexpr (PW == 32) tc::tmem_ld32(t, o); else if constexpr (PW == 16) tc::tmem_ld16(t, o); else tc::tmem_ld8(t, o);
            };
            ld(0, v);
#pragma unroll
            for (int acc = 1; acc < NACC; ++acc) {           // fixed order: deterministic
                ld(acc, u);
#pragma unroll
                for (int i = 0; i < PW; ++i) v[i] += u[i];
            }
            float4* dst = reinterpret_cast<float4*>(xbuf + (size_t)row * XP + p * PW);
#pragma unroll
            for (int c4 = 0; c4 < PW / 4; ++c4) dst[c4] = make_float4(v[4 * c4], v[4 * c4 + 1], v[4 * c4 + 2], v[4 * c4 + 3]);
        }
        asm volatile("bar.sync 2, %0;" ::"n"(WORK_WARPS * 32) : "memory");
        constexpr int C4 = BN / 4;
        for (int c = wt; c < BM * C4; c += WORK_WARPS * 32) {
            const int r = c / C4, c4 = c - r * C4;
            const float4 t = *reinterpret_cast<const float4*>(xbuf + (size_t)r * XP + c4 * 4);
            const float acc[4] = {t.x, t.y, t.z, t.w};
            const int n = n0 + c4 * 4;
            if (n < N) epi.template apply<4>(m0 + r, n, acc);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(NACC * BN) : "memory");
    }
}

// 2-D fp32 tensor map: `inner` contiguous elements per row, `outer` rows, row pitch ld elements; box = 32 x box_outer.
inline int make_map_f32(CUtensorMap* map, const void* ptr, uint64_t inner, uint64_t outer, int64_t ld, uint32_t box_outer) {
    tc::EncodeTiledFn fn = tc::get_encode_fn();
    if (!fn) { sat_set_error("cuTensorMapEncodeTiled entry point not found"); return SAT_ERR_CUDA; }
    cuuint64_t dims[2] = {inner, outer};
    cuuint64_t strides[1] = {(cuuint64_t)ld * 4};
    cuuint32_t box[2] = {(cuuint32_t)BK, box_outer};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { sat_set_error("cuTensorMapEncodeTiled (fp32) failed: %d", (int)r); return SAT_ERR_CUDA; }
    return SAT_OK;
}

// does this GEMM qualify?  both operands K-major, TMA-compatible pitches, and enough work to be worth a 128-row tile
inline bool applies(const float* A, int64_t lda, bool a_k, const float* B, int64_t ldb, bool b_k, int M, int N, int K) {
    return g_fp32_gemm_mode == 1 && a_k && b_k && M >= 16 && N >= 32 && K >= 32 && (lda % 4 == 0) && (ldb % 4 == 0) &&
           (((uintptr_t)A | (uintptr_t)B) & 15) == 0;
}

template <int BN, typename Epi, bool BSPLIT>
int launch_bn(cudaStream_t stream, const CUtensorMap& mA, const CUtensorMap& mB, const CUtensorMap& mBl, int M, int N, int K, const Epi& epi) {
    auto kern = gemm_kernel<BN, Epi, BSPLIT>;
    constexpr int smem = Cfg<BN>::SMEM_BYTES;
    static SatPerDevice configured;
    SAT_TRY(configured.once([&]() -> int {
        SAT_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        return SAT_OK;
    }));
    cudaLaunchConfig_t cfg;
    cudaLaunchAttribute attr[1];
    pdl_config(&cfg, attr, dim3(sat_cdiv(N, BN), sat_cdiv(M, BM)), dim3(THREADS), smem, stream);
    SAT_CHECK_CUDA(cudaLaunchKernelEx(&cfg, kern, mA, mB, mBl, M, N, K, epi));
    return SAT_OK;
}

// B_lo != NULL: B holds tf32(B) and B_lo the remainder (same layout), see split_operand()
template <typename Epi>
int launch(cudaStream_t stream, const float* A, int64_t lda, const float* B, int64_t ldb, int M, int N, int K, const Epi& epi,
           const float* B_lo = nullptr) {
    if (M <= 0 || N <= 0) return SAT_OK;
    // widest tile that keeps the grid in the fewest waves; a k-block costs ~0.41 us of MMAs at BN = 128 (12 x 128x128x8 tf32),
    // half / a quarter of that at 64 / 32 with a ~0.16 us floor (TMA + split), plus ~6 us of launch, prologue and drain
    static int sms = 0;
    if (!sms) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) sms = 148;
    }
    const int nk = sat_cdiv(K, BK), mt = sat_cdiv(M, BM);
    const int cands[3] = {128, 64, 32};
    int bn = 32;
    double best = 1e30;
    for (int i = 0; i < 3; ++i) {
        const int b = cands[i];
        if (b > 32 && b / 2 >= N) continue;
        const long long tiles = (long long)mt * sat_cdiv(N, b);
        const double kbc = 0.41 * b / 128.0 > 0.16 ? 0.41 * b / 128.0 : 0.16;
        const double cost = (double)((tiles + sms - 1) / sms) * (6.0 + nk * kbc);
        if (cost < best * 0.999) { best = cost; bn = b; }
    }
    CUtensorMap mA, mB;
    SAT_TRY(make_map_f32(&mA, A, (uint64_t)K, (uint64_t)M, lda, BM));
    SAT_TRY(make_map_f32(&mB, B, (uint64_t)K, (uint64_t)N, ldb, (uint32_t)bn));
    if (B_lo) {
        CUtensorMap mBl;
        SAT_TRY(make_map_f32(&mBl, B_lo, (uint64_t)K, (uint64_t)N, ldb, (uint32_t)bn));
        if (bn == 128) return launch_bn<128, Epi, true>(stream, mA, mB, mBl, M, N, K, epi);
        if (bn == 64) return launch_bn<64, Epi, true>(stream, mA, mB, mBl, M, N, K, epi);
        return launch_bn<32, Epi, true>(stream, mA, mB, mBl, M, N, K, epi);
    }
    if (bn == 128) return launch_bn<128, Epi, false>(stream, mA, mB, mB, M, N, K, epi);
    if (bn == 64) return launch_bn<64, Epi, false>(stream, mA, mB, mB, M, N, K, epi);
    return launch_bn<32, Epi, false>(stream, mA, mB, mB, M, N, K, epi);
}

// hi = tf32(src) (round to nearest, ties away), lo = src - hi: the operand split of 3xTF32, done once for an operand many launches share
__global__ void split_operand_kernel(float* __restrict__ hi, float* __restrict__ lo, const float* __restrict__ src, int64_t n4) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
        const float4 x = reinterpret_cast<const float4*>(src)[i];
        float4 h, l;
        h.x = tf32_rna(x.x); h.y = tf32_rna(x.y); h.z = tf32_rna(x.z); h.w = tf32_rna(x.w);
        l.x = x.x - h.x; l.y = x.y - h.y; l.z = x.z - h.z; l.w = x.w - h.w;
        reinterpret_cast<float4*>(hi)[i] = h;
        reinterpret_cast<float4*>(lo)[i] = l;
    }
}
inline int split_operand(cudaStream_t stream, float* hi, float* lo, const float* src, int64_t n) {
    SAT_REQUIRE(n % 4 == 0 && ((((uintptr_t)hi | (uintptr_t)lo | (uintptr_t)src) & 15) == 0), "split_operand: 16-byte alignment");
    const int64_t n4 = n / 4;
    int64_t blocks = (n4 + 255) / 256;
    if (blocks > 4 * 148) blocks = 4 * 148;
    if (blocks < 1) return SAT_OK;
    split_operand_kernel<<<(int)blocks, 256, 0, stream>>>(hi, lo, src, n4);
    SAT_CHECK_CUDA(cudaGetLastError());
    return SAT_OK;
}
}  // namespace tc32
