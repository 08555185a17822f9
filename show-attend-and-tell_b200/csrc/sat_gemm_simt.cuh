// True-fp32-accumulate SIMT GEMM:  C[M,N] = A[M,K] * B[N,K]^T  with a fused epilogue functor.
// Used for the fp32 compute mode (config c1 / bit-exact beam search c4, where tensor-core input
// rounding is not allowed) and, with T=bf16, as a bring-up cross-check of the tcgen05 kernel.
// A is "K-major" when A[m*lda + k], else "MN-major" A[k*lda + m]; same for B with n.
#pragma once
#include "sat_common.cuh"

namespace simt {
constexpr int BM = 64, BN = 64, BK = 16, THREADS = 256, PAD = 4;

template <typename T> __device__ __forceinline__ void load4(const T* p, bool vec, int valid, int64_t stride, float* out);

// loads 4 logically consecutive elements (stride 1) starting at p, `valid` of them in range
template <typename T>
__device__ __forceinline__ void load4_contig(const T* p, bool vec, int valid, float* out) {
    if (vec && valid == 4) {
        if (sizeof(T) == 4) {
            float4 v = *reinterpret_cast<const float4*>(p);
            out[0] = v.x; out[1] = v.y; out[2] = v.z; out[3] = v.w;
        } else {
            uint2 u = *reinterpret_cast<const uint2*>(p);
            out[0] = __uint_as_float(u.x << 16); out[1] = __uint_as_float(u.x & 0xffff0000u);
            out[2] = __uint_as_float(u.y << 16); out[3] = __uint_as_float(u.y & 0xffff0000u);
        }
    } else {
#pragma unroll
        for (int i = 0; i < 4; ++i) out[i] = (i < valid) ? to_f32<T>(p[i]) : 0.0f;
    }
}

template <typename T, bool A_KMAJOR, bool B_KMAJOR, typename Epi>
__global__ void __launch_bounds__(THREADS)
gemm_kernel(const T* __restrict__ A, int64_t lda, const T* __restrict__ B, int64_t ldb,
            int M, int N, int K, int vecA, int vecB, Epi epi) {
    __shared__ __align__(16) float As[2][BK][BM + PAD];
    __shared__ __align__(16) float Bs[2][BK][BN + PAD];
    const int tid = threadIdx.x;
    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
    const int ty = tid / 16, tx = tid % 16;

    float ra[4], rb[4];
    auto gload = [&](int k0) {
        if (A_KMAJOR) {
            const int r = tid / 4, kk = (tid % 4) * 4;
            const int m = m0 + r, k = k0 + kk;
            int valid = (m < M) ? max(0, min(4, K - k)) : 0;
            load4_contig<T>(A + (int64_t)m * lda + k, vecA, valid, ra);
        } else {
            const int kk = tid / 16, mm = (tid % 16) * 4;
            const int k = k0 + kk, m = m0 + mm;
            int valid = (k < K) ? max(0, min(4, M - m)) : 0;
            load4_contig<T>(A + (int64_t)k * lda + m, vecA, valid, ra);
        }
        if (B_KMAJOR) {
            const int r = tid / 4, kk = (tid % 4) * 4;
            const int n = n0 + r, k = k0 + kk;
            int valid = (n < N) ? max(0, min(4, K - k)) : 0;
            load4_contig<T>(B + (int64_t)n * ldb + k, vecB, valid, rb);
        } else {
            const int kk = tid / 16, nn = (tid % 16) * 4;
            const int k = k0 + kk, n = n0 + nn;
            int valid = (k < K) ? max(0, min(4, N - n)) : 0;
            load4_contig<T>(B + (int64_t)k * ldb + n, vecB, valid, rb);
        }
    };
    auto sstore = [&](int buf) {
        if (A_KMAJOR) {
            const int r = tid / 4, kk = (tid % 4) * 4;
#pragma unroll
            for (int i = 0; i < 4; ++i) As[buf][kk + i][r] = ra[i];
        } else {
            const int kk = tid / 16, mm = (tid % 16) * 4;
            *reinterpret_cast<float4*>(&As[buf][kk][mm]) = make_float4(ra[0], ra[1], ra[2], ra[3]);
        }
        if (B_KMAJOR) {
            const int r = tid / 4, kk = (tid % 4) * 4;
#pragma unroll
            for (int i = 0; i < 4; ++i) Bs[buf][kk + i][r] = rb[i];
        } else {
            const int kk = tid / 16, nn = (tid % 16) * 4;
            *reinterpret_cast<float4*>(&Bs[buf][kk][nn]) = make_float4(rb[0], rb[1], rb[2], rb[3]);
        }
    };

    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.0f;

    const int nk = (K + BK - 1) / BK;
    gload(0);
    sstore(0);
    __syncthreads();
    for (int kb = 0; kb < nk; ++kb) {
        const int buf = kb & 1;
        if (kb + 1 < nk) gload((kb + 1) * BK);
#pragma unroll
        for (int k = 0; k < BK; ++k) {
            const float4 a4 = *reinterpret_cast<const float4*>(&As[buf][k][ty * 4]);
            const float4 b4 = *reinterpret_cast<const float4*>(&Bs[buf][k][tx * 4]);
            const float a[4] = {a4.x, a4.y, a4.z, a4.w};
            const float b[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
        }
        if (kb + 1 < nk) sstore(buf ^ 1);
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) epi.template apply<4>(m0 + ty * 4 + i, n0 + tx * 4, acc[i]);
}

template <typename T, typename Epi>
int launch(cudaStream_t stream, const T* A, int64_t lda, bool a_kmajor, const T* B, int64_t ldb, bool b_kmajor,
           int M, int N, int K, const Epi& epi) {
    if (M <= 0 || N <= 0) return SAT_OK;
    dim3 grid(sat_cdiv(N, BN), sat_cdiv(M, BM));
    const int epv = 16 / (int)sizeof(T) >= 8 ? 4 : 4;  // 4 elements per load either way
    (void)epv;
    const size_t al = sizeof(T) * 4;
    const int vecA = ((uintptr_t)A % al == 0) && (lda % 4 == 0);
    const int vecB = ((uintptr_t)B % al == 0) && (ldb % 4 == 0);
#define SAT_SIMT_LAUNCH(AK, BK_)                                                                          \
    gemm_kernel<T, AK, BK_, Epi><<<grid, THREADS, 0, stream>>>(A, lda, B, ldb, M, N, K, vecA, vecB, epi)
    if (a_kmajor && b_kmajor) SAT_SIMT_LAUNCH(true, true);
    else if (a_kmajor && !b_kmajor) SAT_SIMT_LAUNCH(true, false);
    else if (!a_kmajor && b_kmajor) SAT_SIMT_LAUNCH(false, true);
    else SAT_SIMT_LAUNCH(false, false);
#undef SAT_SIMT_LAUNCH
    SAT_CHECK_CUDA(cudaGetLastError());
    return SAT_OK;
}
}  // namespace simt
