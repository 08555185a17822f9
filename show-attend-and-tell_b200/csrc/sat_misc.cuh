// Small data-movement / pointwise kernels around the three hot kernels (GEMM, attention step, LSTM epilogue).
#pragma once
#include "sat_common.cuh"

namespace misc {

// dst[perm(r), c] = (TO) src[r, c0 + c]; perm interleaves the four LSTM gate blocks when ilH > 0:
// source row g*ilH + j  ->  destination row 4*j + g.
template <typename TO>
__global__ void pack_weight_kernel(TO* __restrict__ dst, int64_t ld_dst, const float* __restrict__ src, int64_t ld_src,
                                   int rows, int cols, int c0, int ilH) {
    const int r = blockIdx.y;
    const int rd = ilH > 0 ? ((r % ilH) * 4 + r / ilH) : r;
    for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < cols; c += gridDim.x * blockDim.x)
        dst[(int64_t)rd * ld_dst + c] = from_f32<TO>(src[(int64_t)r * ld_src + c0 + c]);
}
template <typename TO>
int pack_weight(cudaStream_t s, TO* dst, int64_t ld_dst, const float* src, int64_t ld_src, int rows, int cols, int c0, int ilH) {
    if (rows <= 0 || cols <= 0) return SAT_OK;
    dim3 grid(min(sat_cdiv(cols, 256), 8), rows);
    pack_weight_kernel<TO><<<grid, 256, 0, s>>>(dst, ld_dst, src, ld_src, rows, cols, c0, ilH);
    SAT_CHECK_CUDA(cudaGetLastError());
    return SAT_OK;
}
// dst[perm(i)] = a[i] + (b ? b[i] : 0)
__global__ void pack_bias_kernel(float* dst, const float* a, const float* b, int n, int ilH) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int d = ilH > 0 ? ((i % ilH) * 4 + i / ilH) : i;
    dst[d] = a[i] + (b ? b[i] : 0.0f);
}
inline int pack_bias(cudaStream_t s, float* dst, const float* a, const float* b, int n, int ilH) {
    pack_bias_kernel<<<sat_cdiv(n, 256), 256, 0, s>>>(dst, a, b, n, ilH);
    SAT_CHECK_CUDA(cudaGetLastError());
    return SAT_OK;
}
// All of a call's weight / bias packing in ONE launch: the jobs travel as a kernel parameter table, one block per row
// (weights) or per 256 elements (biases).  16 separate 2-6 us launches cost ~95 us of mostly launch gaps per train step.
struct PackJob {
    void* dst; const float* src; const float* src2;
    int64_t ld_dst, ld_src;
    int rows, cols, c0, ilH;       // bias job: rows = blocks of 256 elements, cols = n
    int first_block, is_bias;
};
constexpr int PACK_MAX_JOBS = 24;
struct PackTable { PackJob job[PACK_MAX_JOBS]; int njobs; };
template <typename TO>
__global__ void __launch_bounds__(256) pack_all_kernel(const __grid_constant__ PackTable tab) {
    int j = 0;
    while (j + 1 < tab.njobs && (int)blockIdx.x >= tab.job[j + 1].first_block) ++j;
    const PackJob& J = tab.job[j];
    const int r = blockIdx.x - J.first_block;
    if (J.is_bias) {
        const int i = r * 256 + threadIdx.x;
        if (i >= J.cols) return;
        const int d = J.ilH > 0 ? ((i % J.ilH) * 4 + i / J.ilH) : i;
        reinterpret_cast<float*>(J.dst)[d] = J.src[i] + (J.src2 ? J.src2[i] : 0.0f);
        return;
    }
    const int rd = J.ilH > 0 ? ((r % J.ilH) * 4 + r / J.ilH) : r;
    TO* dst = reinterpret_cast<TO*>(J.dst) + (int64_t)rd * J.ld_dst;
    const float* src = J.src + (int64_t)r * J.ld_src + J.c0;
    if ((J.cols & 3) == 0 && (((uintptr_t)src | (uintptr_t)dst) & 15) == 0 && (sizeof(TO) == 4 || ((uintptr_t)dst & 7) == 0)) {
        for (int c = 4 * threadIdx.x; c < J.cols; c += 1024) {       // 16-byte loads, 8- / 16-byte stores
            const float4 v = __ldg(reinterpret_cast<const float4*>(src + c));
            const float f[4] = {v.x, v.y, v.z, v.w};
            store4<TO>(dst + c, f);
        }
    } else {
        for (int c = threadIdx.x; c < J.cols; c += 256) dst[c] = from_f32<TO>(src[c]);
    }
}
template <typename TO>
struct PackList {
    PackTable tab;
    int blocks;
    PackList() : blocks(0) { tab.njobs = 0; }
    int push(void* dst, int64_t ld_dst, const float* src, const float* src2, int64_t ld_src, int rows, int cols, int c0, int ilH, int is_bias) {
        if (rows <= 0 || cols <= 0) return SAT_OK;
        SAT_REQUIRE(tab.njobs < PACK_MAX_JOBS, "too many pack jobs");
        PackJob& J = tab.job[tab.njobs++];
        J.dst = dst; J.src = src; J.src2 = src2; J.ld_dst = ld_dst; J.ld_src = ld_src;
        J.rows = rows; J.cols = cols; J.c0 = c0; J.ilH = ilH; J.first_block = blocks; J.is_bias = is_bias;
        blocks += rows;
        return SAT_OK;
    }
    // same arguments as pack_weight / pack_bias
    int weight(TO* dst, int64_t ld_dst, const float* src, int64_t ld_src, int rows, int cols, int c0, int ilH) {
        return push(dst, ld_dst, src, nullptr, ld_src, rows, cols, c0, ilH, 0);
    }
    int bias(float* dst, const float* a, const float* b, int n, int ilH) {
        return push(dst, 0, a, b, 0, sat_cdiv(n, 256), n, 0, ilH, 1);
    }
    int flush(cudaStream_t s) {
        if (tab.njobs == 0) return SAT_OK;
        pack_all_kernel<TO><<<blocks, 256, 0, s>>>(tab);
        SAT_CHECK_CUDA(cudaGetLastError());
        tab.njobs = 0; blocks = 0;
        return SAT_OK;
    }
};
// rank[r] = position of the target's logit in row r's descending order (utils.py:44-80 sequence_accuracy asks whether the
// target is among the top k = whether rank < k; one pass over the logits instead of two topk passes).  Ties are broken
// towards the LOWER index, so at most k entries of a tied block are admitted, like `preds.topk(k)` (which entries of a tie
// topk returns is unspecified upstream; in ado mode the final ReLU, decoder.py:158, makes blocks of exact zeros the norm).
// A target outside [0, V) or a NaN target logit gets rank V (never correct).  One block per row.
__global__ void __launch_bounds__(256) logit_rank_kernel(const float* __restrict__ logits, int64_t ld, int V,
                                                         const int64_t* __restrict__ targets, int32_t* __restrict__ rank) {
    __shared__ int red[8];
    const int64_t r = blockIdx.x;
    const int64_t tgt = targets[r];
    if (tgt < 0 || tgt >= V) { if (threadIdx.x == 0) rank[r] = V; return; }
    const float* p = logits + r * ld;
    const float x = p[tgt];
    if (x != x) { if (threadIdx.x == 0) rank[r] = V; return; }
    int n = 0;
    for (int v = threadIdx.x; v < V; v += blockDim.x) {
        const float y = p[v];
        n += (y > x || (y == x && v < (int)tgt) || y != y) ? 1 : 0;       // a NaN logit sorts above everything (torch.topk)
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) n += __shfl_xor_sync(0xffffffffu, n, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = n;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
        for (int w = 0; w < 8; ++w) t += red[w];
        rank[r] = t;
    }
}

// utils.py:101-107 calculate_caption_lengths: number of tokens not in {skip0, skip1, skip2} (train.py:174-177 passes pad / start /
// eos).  One block, fixed reduction order; the count stays on the device (the reference does .item() every batch).
__global__ void __launch_bounds__(1024) count_tokens_kernel(const int64_t* __restrict__ tok, int64_t rows, int64_t cols, int64_t stride,
                                                            int64_t s0, int64_t s1, int64_t s2, int64_t* __restrict__ out) {
    __shared__ int red[32];
    int n = 0;
    const int64_t total = rows * cols;
    for (int64_t i = threadIdx.x; i < total; i += blockDim.x) {
        const int64_t r = i / cols, c = i - r * cols;
        const int64_t t = tok[r * stride + c];
        n += (t != s0 && t != s1 && t != s2) ? 1 : 0;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) n += __shfl_xor_sync(0xffffffffu, n, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = n;
    __syncthreads();
    if (threadIdx.x == 0) {
        int64_t t = 0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += red[w];
        out[0] = t;
    }
}

// N2 feature cache: out[i, :] = store[idx[i], :] for rows of `row_vec16` 16-byte vectors (frozen-encoder features kept in HBM,
// encoder.py:29-31 / train.py:132).  An index outside [0, n_items) gives a zero row and sets *err_flag.
__global__ void __launch_bounds__(256) gather_items_kernel(const uint4* __restrict__ store, int64_t row_vec16, int64_t n_items,
                                                           const int64_t* __restrict__ idx, uint4* __restrict__ out, int* __restrict__ err_flag) {
    const int64_t i = blockIdx.y;
    const int64_t src = idx[i];
    const bool ok = (src >= 0 && src < n_items);
    if (!ok && err_flag && threadIdx.x == 0 && blockIdx.x == 0) atomicOr(err_flag, 1);
    const uint4* s = store + src * row_vec16;
    uint4* d = out + i * row_vec16;
    for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < row_vec16; v += (int64_t)gridDim.x * blockDim.x) {
        uint4 x = make_uint4(0u, 0u, 0u, 0u);
        if (ok) x = ldg_stream16(s + v);
        asm volatile("st.global.cs.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(d + v), "r"(x.x), "r"(x.y), "r"(x.z), "r"(x.w) : "memory");
    }
}

// inverse of pack_weight for gradients: dst_fp32[r, c0 + c] = src[perm(r), c]   (dst row pitch ld_dst)
__global__ void unpack_grad_kernel(float* __restrict__ dst, int64_t ld_dst, const float* __restrict__ src, int64_t ld_src,
                                   int rows, int cols, int c0, int ilH) {
    const int r = blockIdx.y;
    const int rs = ilH > 0 ? ((r % ilH) * 4 + r / ilH) : r;
    for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < cols; c += gridDim.x * blockDim.x)
        dst[(int64_t)r * ld_dst + c0 + c] = src[(int64_t)rs * ld_src + c];
}
inline int unpack_grad(cudaStream_t s, float* dst, int64_t ld_dst, const float* src, int64_t ld_src, int rows, int cols, int c0, int ilH) {
    if (rows <= 0 || cols <= 0) return SAT_OK;
    dim3 grid(min(sat_cdiv(cols, 256), 8), rows);
    unpack_grad_kernel<<<grid, 256, 0, s>>>(dst, ld_dst, src, ld_src, rows, cols, c0, ilH);
    SAT_CHECK_CUDA(cudaGetLastError());
    return SAT_OK;
}

// strided (B,L,D) of TI -> contiguous (B,L,D) of TO   (Q20: encoder output is a permuted view)
template <typename TI, typename TO>
__global__ void cast3d_kernel(TO* __restrict__ dst, const TI* __restrict__ src, int B, int L, int D,
                              int64_t sb, int64_t sl, int64_t sd) {
    __shared__ float tile[32][33];
    // tile over (l, d) so both the strided read (l fastest when sd > sl) and the write (d fastest) coalesce
    const int b = blockIdx.z;
    const int l0 = blockIdx.y * 32, d0 = blockIdx.x * 32;
    const bool l_fast = (sl < sd);
    if (l_fast) {
        for (int i = threadIdx.y; i < 32; i += blockDim.y) {
            const int d = d0 + i, l = l0 + threadIdx.x;
            if (d < D && l < L) tile[i][threadIdx.x] = to_f32<TI>(src[b * sb + l * sl + d * sd]);
        }
        __syncthreads();
        for (int i = threadIdx.y; i < 32; i += blockDim.y) {
            const int l = l0 + i, d = d0 + threadIdx.x;
            if (l < L && d < D) dst[((int64_t)b * L + l) * D + d] = from_f32<TO>(tile[threadIdx.x][i]);
        }
    } else {
        for (int i = threadIdx.y; i < 32; i += blockDim.y) {
            const int l = l0 + i, d = d0 + threadIdx.x;
            if (l < L && d < D) dst[((int64_t)b * L + l) * D + d] = from_f32<TO>(to_f32<TI>(src[b * sb + l * sl + d * sd]));
        }
    }
}
template <typename TI, typename TO>
int cast3d(cudaStream_t s, TO* dst, const TI* src, int B, int L, int D, int64_t sb, int64_t sl, int64_t sd) {
    dim3 grid(sat_cdiv(D, 32), sat_cdiv(L, 32), B), block(32, 8);
    cast3d_kernel<TI, TO><<<grid, block, 0, s>>>(dst, src, B, L, D, sb, sl, sd);
    SAT_CHECK_CUDA(cudaGetLastError());
    return SAT_OK;
}

// flat cast / scale
template <typename TI, typename TO>
__global__ void cast_kernel(TO* __restrict__ dst, const TI* __restrict__ src, int64_t n, float scale) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        dst[i] = from_f32<TO>(to_f32<TI>(src[i]) * scale);
}
template <typename TI, typename TO>
int cast(cudaStream_t s, TO* dst, const TI* src, int64_t n, float scale = 1.0f) {
    if (n <= 0) return SAT_OK;
    const int blocks = (int)(sat_cdiv64(n, 256) < (int64_t)(148 * 16) ? sat_cdiv64(n, 256) : (int64_t)(148 * 16));
    cast_kernel<TI, TO><<<blocks, 256, 0, s>>>(dst, src, n, scale);
    SAT_CHECK_CUDA(cudaGetLastError());
    return SAT_OK;
}

// out[i] = sum_t src[t*n + i]  (fp32 accumulation; 16-byte vectors; n % (16/sizeof(T)) == 0)
template <typename T>
__global__ void sum_over_t_kernel(T* __restrict__ out, const T* __restrict__ src, int64_t nvec, int64_t n, int Tn) {
    constexpr int VEC = Vec16<T>::N;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
        float acc[VEC];
#pragma unroll
        for (int k = 0; k < VEC; ++k) acc[k] = 0.0f;
#pragma unroll 4
        for (int t = 0; t < Tn; ++t) {
            float f[VEC];
            Vec16<T>::unpack(ldg_stream16(src + (size_t)t * n + i * VEC), f);
#pragma unroll
            for (int k = 0; k < VEC; ++k) acc[k] += f[k];
        }
        *reinterpret_cast<uint4*>(out + i * VEC) = Vec16<T>::pack(acc);
    }
}
template <typename T>
int sum_over_t(cudaStream_t s, T* out, const T* src, int64_t n, int Tn) {
    constexpr int VEC = Vec16<T>::N;
    SAT_REQUIRE(n % VEC == 0, "sum_over_t: n must be a multiple of %d", VEC);
    const int64_t nvec = n / VEC;
    const int blocks = (int)(sat_cdiv64(nvec, 256) < (int64_t)(148 * 8) ? sat_cdiv64(nvec, 256) : (int64_t)(148 * 8));
    sum_over_t_kernel<T><<<blocks, 256, 0, s>>>(out, src, nvec, n, Tn);
    SAT_CHECK_CUDA(cudaGetLastError());
    return SAT_OK;
}

// The same pass with the column sums of the result as a side output (the bias gradient of the encoder projection):
// rows of H columns; a block is `rpb` whole rows of 16-byte vectors wide, so with a grid stride that is a multiple of the
// row length a thread meets the same columns every iteration and keeps their sums in registers.
// part[block][H] holds per-block partial column sums (fp32, taken before rounding to T); colsum_final_kernel adds them.
// SHARE: the launch leaves room on every SM for a tcgen05 GEMM CTA of the other stream (19 warps x 80 registers).  The register
// file is split per SM sub-partition (16 K registers each): the GEMM's 5 warps there take 12 800, so this kernel's 2 warps per
// sub-partition must stay at <= 48 registers (2 x 1536 = 3072; at 64 registers the 4096 did NOT fit and the GEMM CTA was never
// co-scheduled, profiles/r1_tail_overlap.txt).  One block of 256 threads per SM, six loads in flight per thread.
template <typename T, bool SHARE>
__global__ void __launch_bounds__(256, SHARE ? 5 : 4)
sum_over_t_colsum_kernel(T* __restrict__ out, const T* __restrict__ src, int64_t nvec, int64_t n, int Tn, int H,
                         float* __restrict__ part) {
    constexpr int VEC = Vec16<T>::N;
    extern __shared__ float red[];                 // [blockDim][VEC]
    const int cpr = H / VEC;                       // vectors per row
    float cs[VEC];
#pragma unroll
    for (int k = 0; k < VEC; ++k) cs[k] = 0.0f;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
        float acc[VEC];
#pragma unroll
        for (int k = 0; k < VEC; ++k) acc[k] = 0.0f;
        constexpr int U = SHARE ? 6 : 5;
        int t = 0;
        for (; t + U <= Tn; t += U) {
            uint4 raw[U];
#pragma unroll
            for (int u = 0; u < U; ++u) raw[u] = ldg_stream16(src + (size_t)(t + u) * n + i * VEC);
#pragma unroll
            for (int u = 0; u < U; ++u) {
                float f[VEC];
                Vec16<T>::unpack(raw[u], f);
#pragma unroll
                for (int k = 0; k < VEC; ++k) acc[k] += f[k];
            }
        }
        for (; t < Tn; ++t) {
            float f[VEC];
            Vec16<T>::unpack(ldg_stream16(src + (size_t)t * n + i * VEC), f);
#pragma unroll
            for (int k = 0; k < VEC; ++k) acc[k] += f[k];
        }
        *reinterpret_cast<uint4*>(out + i * VEC) = Vec16<T>::pack(acc);
#pragma unroll
        for (int k = 0; k < VEC; ++k) cs[k] += acc[k];
    }
#pragma unroll
    for (int k = 0; k < VEC; ++k) red[threadIdx.x * VEC + k] = cs[k];
    __syncthreads();
    if ((int)threadIdx.x < cpr) {
        const int rpb = blockDim.x / cpr;
#pragma unroll
        for (int k = 0; k < VEC; ++k) {
            float t = 0.0f;
            for (int r = 0; r < rpb; ++r) t += red[(r * cpr + threadIdx.x) * VEC + k];
            part[(size_t)blockIdx.x * H + threadIdx.x * VEC + k] = t;
        }
    }
}
// The same pass built to run BESIDE a tcgen05 GEMM CTA of another stream (backward tail, sat_decoder.cuh): the bytes in flight
// live in shared memory instead of registers.  One 256-thread block per SM (<= 48 registers: the GEMM leaves 3.5 K registers per
// SM sub-partition), two cp.async stages of Tn x CV vectors (~26 KB, what a 193 KB GEMM CTA leaves free).  A chunk is CV
// consecutive 16-byte vectors, shorter than a row, so its vectors are distinct columns: column sums accumulate in shared memory.
template <typename T>
__global__ void __launch_bounds__(256, 5)
sum_over_t_async_kernel(T* __restrict__ out, const T* __restrict__ src, int64_t nvec, int64_t n, int Tn, int H, int CV,
                        float* __restrict__ part) {
    constexpr int VEC = Vec16<T>::N;
    extern __shared__ __align__(16) uint8_t sm_raw[];
    uint4* stage = reinterpret_cast<uint4*>(sm_raw);                       // [2][Tn][CV]
    float* cs = reinterpret_cast<float*>(sm_raw + (size_t)2 * Tn * CV * 16);   // [H]
    const int cpr = H / VEC;
    for (int i = threadIdx.x; i < H; i += blockDim.x) cs[i] = 0.0f;
    const int64_t nchunks = (nvec + CV - 1) / CV;
    auto issue = [&](int64_t k, int s) {
        if (k < nchunks) {
            const int64_t v0 = k * CV;
            const int valid = (int)((nvec - v0) < CV ? (nvec - v0) : CV);
            for (int e = threadIdx.x; e < Tn * CV; e += blockDim.x) {
                const int t = e / CV, v = e - t * CV;
                if (v < valid) {
                    const uint32_t dst = smem_u32(stage + ((size_t)s * Tn + t) * CV + v);
                    const T* g = src + (size_t)t * n + (v0 + v) * VEC;
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(g) : "memory");
                }
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    int64_t k = blockIdx.x;
    issue(k, 0);
    issue(k + gridDim.x, 1);
    for (int it = 0; k < nchunks; k += gridDim.x, ++it) {
        const int s = it & 1;
        asm volatile("cp.async.wait_group 1;" ::: "memory");
        __syncthreads();
        const int64_t v0 = k * CV;
        const int valid = (int)((nvec - v0) < CV ? (nvec - v0) : CV);
        if ((int)threadIdx.x < valid) {
            float acc[VEC];
#pragma unroll
            for (int j = 0; j < VEC; ++j) acc[j] = 0.0f;
            for (int t = 0; t < Tn; ++t) {
                float f[VEC];
                Vec16<T>::unpack(stage[((size_t)s * Tn + t) * CV + threadIdx.x], f);
#pragma unroll
                for (int j = 0; j < VEC; ++j) acc[j] += f[j];
            }
            const int64_t vi = v0 + threadIdx.x;
            *reinterpret_cast<uint4*>(out + vi * VEC) = Vec16<T>::pack(acc);
            float* c = cs + (int)(vi % cpr) * VEC;
#pragma unroll
            for (int j = 0; j < VEC; ++j) c[j] += acc[j];
        }
        __syncthreads();
        issue(k + 2 * (int64_t)gridDim.x, s);
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
    for (int i = threadIdx.x; i < H; i += blockDim.x) part[(size_t)blockIdx.x * H + i] = cs[i];
}
__global__ void colsum_final_kernel(const float* __restrict__ part, int S, int N, float* __restrict__ out, int ilH);
// out (rows x H, T) = sum_t src[t];  colsum_out[H] = column sums of that (fp32)
template <typename T>
int sum_over_t_colsum(cudaStream_t s, T* out, const T* src, int64_t rows, int H, int Tn, float* colsum_out, float* scratch,
                      int64_t scratch_floats, bool share_sms = false) {
    constexpr int VEC = Vec16<T>::N;
    SAT_REQUIRE(H % VEC == 0 && H / VEC <= 1024, "sum_over_t_colsum: H must be a multiple of %d", VEC);
    const int cpr = H / VEC;
    const int rpb = cpr >= 256 ? 1 : 256 / cpr;
    const int threads = cpr * rpb;
    const int64_t n = rows * H, nvec = n / VEC;
    int64_t blocks = sat_cdiv64(nvec, threads);
    share_sms = share_sms && threads == 256;
    const int64_t cap = share_sms ? 148 : 4 * 148;
    if (blocks > cap) blocks = cap;
    if (blocks * H > scratch_floats) blocks = scratch_floats / H;
    SAT_REQUIRE(blocks >= 1, "sum_over_t_colsum: scratch too small");
    if (share_sms) {
        // chunk length: two stages of Tn x CV vectors in ~25 KB, CV <= 256 threads and shorter than a row
        int CV = (int)(12800 / ((int64_t)Tn * 16));
        if (CV > cpr - 1) CV = cpr - 1;
        if (CV > 256) CV = 256;
        if (CV >= 4) {
            CV &= ~3;
            // an SM cannot change its L1 / shared-memory split while blocks are resident: ask for the all-shared split so that
            // a GEMM CTA (~193 KB of shared memory) can land beside this kernel's block
            static SatPerDevice configured;
            SAT_TRY(configured.once([&]() -> int {
                SAT_CHECK_CUDA(cudaFuncSetAttribute(sum_over_t_async_kernel<T>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                                    cudaSharedmemCarveoutMaxShared));
                return SAT_OK;
            }));
            const size_t smem = (size_t)2 * Tn * CV * 16 + (size_t)H * 4;
            sum_over_t_async_kernel<T><<<(int)blocks, 256, smem, s>>>(out, src, nvec, n, Tn, H, CV, scratch);
        } else {
            sum_over_t_colsum_kernel<T, true><<<(int)blocks, threads, (size_t)threads * VEC * 4, s>>>(out, src, nvec, n, Tn, H, scratch);
        }
    }
    else sum_over_t_colsum_kernel<T, false><<<(int)blocks, threads, (size_t)threads * VEC * 4, s>>>(out, src, nvec, n, Tn, H, scratch);
    colsum_final_kernel<<<sat_cdiv(H, 32), 256, 0, s>>>(scratch, (int)blocks, H, colsum_out, 0);
    SAT_CHECK_CUDA(cudaGetLastError());
    return SAT_OK;
}

// ---- sum over time of the attention pre-activation gradient, RECOMPUTED instead of stored -------------------------------------
// Backward of attention.py:16-18 for the encoder projection W:  dWa[b,l,:] = sum_t dpre_t[b,l,:]  with
//   dpre_t[b,l,h] = de_t[b,l] * v[h] * (1 - tanh^2(Wa[b,l,h] + Uh_t[b,h])).
// The streaming backward kernel has every dpre_t in registers, but writing it out costs B*L*H*sizeof(T) per decode step and the
// sum over t reads all of it back (2 x 1.03 GB at c2).  Everything dpre_t depends on besides Wa is tiny -- the score gradients
// de_t (B*L floats per step) and the saved Uh_t rows (B*H floats per step) -- so the backward steps store only de_t and this kernel
// rebuilds the sum in ONE pass over Wa: per (b, l) row it holds Wa in registers and loops over t with Uh_t / de_t in shared memory.
// One MUFU tanh per (t, b, l, h): 514 M at c2, i.e. ~115 us of the SFU pipe instead of ~2 GB of HBM traffic; the sum is kept in
// fp32 and rounded to T once (the stored slabs were rounded per step).  Measured at c2: 162 us alone (ncu: the SFU pipe is the busiest
// unit at 68 %, mio_throttle the top stall); a variant with the next item's rows prefetched by cp.async was slower (170 us) and was dropped.
// Block = (H/4, 2) threads: threadIdx.x owns 4 adjacent columns, threadIdx.y every second row of the item's DS_LC rows.  Items
// (image, row chunk) are dealt to a persistent grid in contiguous ranges, so a block reloads the Uh table only when the image changes
// and keeps its column sums (the bias gradient) in registers: part[block][H], added up by colsum_final_kernel.
constexpr int DS_LC = 14;            // rows per item (196 = 14 x 14)
constexpr int DS_NLR = DS_LC / 2;    // rows per thread
template <typename T, bool PRECISE>
__global__ void __launch_bounds__(384)
dpre_sum_kernel(T* __restrict__ out, const T* __restrict__ wa, const float* __restrict__ uf, int64_t ld_uf,
                const float* __restrict__ de, const float* __restrict__ v, int B, int L, int H, int Tn, int TC,
                int64_t items, int chunks, float* __restrict__ part) {
    extern __shared__ __align__(16) float ds_sm[];
    float* uh_s = ds_sm;                               // [TC][H]
    float* de_s = ds_sm + (size_t)TC * H;              // [TC][DS_LC]
    const int tx = threadIdx.x, ty = threadIdx.y, h = tx * 4;
    const int nthr = blockDim.x * blockDim.y, tid = ty * blockDim.x + tx;
    const int64_t i0 = items * blockIdx.x / gridDim.x, i1 = items * (blockIdx.x + 1) / gridDim.x;
    const float4 v4 = *reinterpret_cast<const float4*>(v + h);
    const float vv[4] = {v4.x, v4.y, v4.z, v4.w};
    float cs[4] = {0.0f, 0.0f, 0.0f, 0.0f};
    int64_t b_loaded = -1;
    for (int64_t it = i0; it < i1; ++it) {
        const int64_t b = it / chunks;
        const int l0 = (int)(it - b * chunks) * DS_LC;
        float w[DS_NLR][4], acc[DS_NLR][4];
#pragma unroll
        for (int i = 0; i < DS_NLR; ++i) {
            const int l = l0 + ty + 2 * i;
#pragma unroll
            for (int k = 0; k < 4; ++k) { acc[i][k] = 0.0f; w[i][k] = 0.0f; }
            if (l < L) load4<T>(wa + ((size_t)b * L + l) * H + h, w[i]);
        }
        for (int t0 = 0; t0 < Tn; t0 += TC) {
            const int tn = min(TC, Tn - t0);
            const bool reload = (b != b_loaded) || (TC < Tn);
            __syncthreads();                                         // the previous item / chunk is done with the tables
            if (reload)
                for (int i = tid; i < tn * (H / 4); i += nthr) {
                    const int t = i / (H / 4), c = i - t * (H / 4);
                    *reinterpret_cast<float4*>(uh_s + (size_t)t * H + c * 4) =
                        *reinterpret_cast<const float4*>(uf + ((size_t)(t0 + t) * B + b) * ld_uf + c * 4);
                }
            for (int i = tid; i < tn * DS_LC; i += nthr) {
                const int t = i / DS_LC, r = i - t * DS_LC;
                de_s[i] = (l0 + r < L) ? de[((size_t)(t0 + t) * B + b) * L + l0 + r] : 0.0f;
            }
            __syncthreads();
            for (int t = 0; t < tn; ++t) {
                const float4 u4 = *reinterpret_cast<const float4*>(uh_s + (size_t)t * H + h);
                const float u[4] = {u4.x, u4.y, u4.z, u4.w};
#pragma unroll
                for (int i = 0; i < DS_NLR; ++i) {
                    const float d = de_s[t * DS_LC + ty + 2 * i];
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const float th = attn_tanh<PRECISE>(w[i][k] + u[k]);
                        acc[i][k] = fmaf(d, fmaf(-th, th, 1.0f), acc[i][k]);
                    }
                }
            }
        }
        b_loaded = b;
#pragma unroll
        for (int i = 0; i < DS_NLR; ++i) {
            const int l = l0 + ty + 2 * i;
            if (l < L) {
#pragma unroll
                for (int k = 0; k < 4; ++k) { acc[i][k] *= vv[k]; cs[k] += acc[i][k]; }
                store4<T>(out + ((size_t)b * L + l) * H + h, acc[i]);
            }
        }
    }
    if (part) {                                                      // the two row-parities of a block: through shared memory
        __syncthreads();
        if (ty == 1) *reinterpret_cast<float4*>(ds_sm + h) = make_float4(cs[0], cs[1], cs[2], cs[3]);
        __syncthreads();
        if (ty == 0) {
            const float4 o = *reinterpret_cast<const float4*>(ds_sm + h);
            *reinterpret_cast<float4*>(part + (size_t)blockIdx.x * H + h) = make_float4(cs[0] + o.x, cs[1] + o.y, cs[2] + o.z, cs[3] + o.w);
        }
    }
}
// out (B, L, H) T = sum_t dpre_t rebuilt from wa (B, L, H), uf (T, B, ld_uf: first H columns = U h_t + b_U), de (T, B, L) and v (H);
// colsum_out[H] (may be NULL) = its column sums over (b, l), fp32.  scratch: blocks x H floats.
template <typename T, bool PRECISE>
int dpre_sum(cudaStream_t s, T* out, const T* wa, const float* uf, int64_t ld_uf, const float* de, const float* v, int B, int L,
             int H, int Tn, float* colsum_out, float* scratch, int64_t scratch_floats) {
    SAT_REQUIRE(H % 4 == 0 && H / 4 <= 192 && ld_uf % 4 == 0, "dpre_sum: H (%d) must be a multiple of 4, at most 768", H);
    const int chunks = sat_cdiv(L, DS_LC);
    const int64_t items = (int64_t)B * chunks;
    int TC = Tn;
    const int tc_max = (int)((96 * 1024) / ((size_t)(H + DS_LC) * 4));
    if (TC > tc_max) TC = tc_max;
    const size_t smem = (size_t)TC * (H + DS_LC) * 4;
    int64_t blocks = items < 3 * 148 ? items : 3 * 148;
    if (colsum_out && blocks * H > scratch_floats) blocks = scratch_floats / H;
    SAT_REQUIRE(blocks >= 1, "dpre_sum: scratch too small");
    static SatPerDevice configured;
    SAT_TRY(configured.once([&]() -> int {
        SAT_CHECK_CUDA(cudaFuncSetAttribute(dpre_sum_kernel<T, PRECISE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024 + 1024));
        return SAT_OK;
    }));
    dpre_sum_kernel<T, PRECISE><<<(int)blocks, dim3(H / 4, 2), smem, s>>>(out, wa, uf, ld_uf, de, v, B, L, H, Tn, TC, items, chunks,
                                                                           colsum_out ? scratch : nullptr);
    SAT_CHECK_CUDA(cudaGetLastError());
    if (colsum_out) {
        colsum_final_kernel<<<sat_cdiv(H, 32), 256, 0, s>>>(scratch, (int)blocks, H, colsum_out, 0);
        SAT_CHECK_CUDA(cudaGetLastError());
    }
    return SAT_OK;
}

// abar[b,d] = mean_l a[b,l,d]   (decoder.py:139 / :104).  One pass over `a` (205 MB at c2): 16-byte loads, the L loop split
// over blockIdx.z so that enough loads are in flight, partial sums combined with a second tiny pass when split.
template <typename T>
__global__ void mean_l_kernel(const T* __restrict__ a, float* __restrict__ part, int L, int D, int lsplit) {
    constexpr int VEC = Vec16<T>::N;
    const int b = blockIdx.y, z = blockIdx.z;
    const int dv = blockIdx.x * blockDim.x + threadIdx.x;          // 16-byte column group
    if (dv * VEC >= D) return;
    const int l0 = (int)((int64_t)L * z / lsplit), l1 = (int)((int64_t)L * (z + 1) / lsplit);
    const T* p = a + (size_t)b * L * D + (size_t)dv * VEC;
    float acc[VEC];
#pragma unroll
    for (int k = 0; k < VEC; ++k) acc[k] = 0.0f;
#pragma unroll 4
    for (int l = l0; l < l1; ++l) {
        float f[VEC];
        Vec16<T>::unpack(ldg_stream16(p + (size_t)l * D), f);
#pragma unroll
        for (int k = 0; k < VEC; ++k) acc[k] += f[k];
    }
    float* o = part + ((size_t)z * gridDim.y + b) * D + (size_t)dv * VEC;
#pragma unroll
    for (int k = 0; k < VEC; ++k) o[k] = acc[k];
}
template <typename T>
__global__ void mean_l_final_kernel(const float* __restrict__ part, float* __restrict__ out_f, T* __restrict__ out_t, int64_t n, int lsplit, float inv_l) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float s = 0.0f;
    for (int z = 0; z < lsplit; ++z) s += part[(size_t)z * n + i];
    s *= inv_l;
    if (out_f) out_f[i] = s;
    if (out_t) out_t[i] = from_f32<T>(s);
}
// scratch: lsplit * B * D floats
template <typename T>
int mean_l(cudaStream_t s, const T* a, float* out_f, T* out_t, int B, int L, int D, float* scratch, int64_t scratch_floats) {
    constexpr int VEC = Vec16<T>::N;
    SAT_REQUIRE(D % VEC == 0, "mean_l: D must be a multiple of %d", VEC);
    int lsplit = 4;
    while (lsplit > 1 && ((int64_t)lsplit * B * D > scratch_floats || lsplit > L)) lsplit /= 2;
    SAT_REQUIRE((int64_t)lsplit * B * D <= scratch_floats, "mean_l: scratch too small");
    dim3 grid(sat_cdiv(D / VEC, 128), B, lsplit);
    mean_l_kernel<T><<<grid, 128, 0, s>>>(a, scratch, L, D, lsplit);
    const int64_t n = (int64_t)B * D;
    mean_l_final_kernel<T><<<(int)sat_cdiv64(n, 256), 256, 0, s>>>(scratch, out_f, out_t, n, lsplit, 1.0f / (float)L);
    SAT_CHECK_CUDA(cudaGetLastError());
    return SAT_OK;
}

// out[r, 0:H] = (TO) table[idx[r*idx_stride], 0:H]   (embedding gather, decoder.py:87/:89/:133/:183)
template <typename TO>
__global__ void gather_rows_kernel(TO* __restrict__ out, int64_t ld_out, const float* __restrict__ table, int H,
                                   const int64_t* __restrict__ idx, int64_t idx_stride, int V) {
    const int r = blockIdx.x;
    const int64_t tok = idx[(int64_t)r * idx_stride];
    if (tok < 0 || tok >= V) {       // ids produced on the device (argmax / top-k) are always in range; a zero row otherwise
        for (int h = threadIdx.x; h < H; h += blockDim.x) out[(int64_t)r * ld_out + h] = from_f32<TO>(0.0f);
        return;
    }
    const float* src = table + (size_t)tok * H;
    for (int h = threadIdx.x; h < H; h += blockDim.x) out[(int64_t)r * ld_out + h] = from_f32<TO>(src[h]);
}
template <typename TO>
int gather_rows(cudaStream_t s, TO* out, int64_t ld_out, const float* table, int H, const int64_t* idx, int64_t idx_stride, int rows, int V) {
    if (rows <= 0) return SAT_OK;
    gather_rows_kernel<TO><<<rows, 128, 0, s>>>(out, ld_out, table, H, idx, idx_stride, V);
    SAT_CHECK_CUDA(cudaGetLastError());
    return SAT_OK;
}
// 2-D gather for (B, T) token grids written TIME-MAJOR to (T, B, H): row r = t*B + b reads cap[b*cap_stride + t]
// A token outside [0, V) (nn.Embedding raises IndexError upstream) gathers a zero row -- matching the scatter below, which
// skips it -- and sets SAT_FLAG_BAD_TOKEN in *err_flag; the host wrapper turns the flag into an exception.
template <typename TO>
__global__ void gather_bt_kernel(TO* __restrict__ out, const float* __restrict__ table, int H,
                                 const int64_t* __restrict__ cap, int64_t cap_stride, int B, int V, int* __restrict__ err_flag) {
    const int r = blockIdx.x, t = r / B, b = r - t * B;
    const int64_t tok = cap[(int64_t)b * cap_stride + t];
    if (tok < 0 || tok >= V) {
        if (err_flag && threadIdx.x == 0) atomicOr(err_flag, SAT_FLAG_BAD_TOKEN);
        for (int h = threadIdx.x; h < H; h += blockDim.x) out[(int64_t)r * H + h] = from_f32<TO>(0.0f);
        return;
    }
    const float* src = table + (size_t)tok * H;
    for (int h = threadIdx.x; h < H; h += blockDim.x) out[(int64_t)r * H + h] = from_f32<TO>(src[h]);
}
// grad scatter: table_grad[tok, :] += dE[r, :], dE time-major (row r = t*B + b)
template <typename TI>
__global__ void scatter_add_bt_kernel(float* __restrict__ table_grad, int H, const TI* __restrict__ dE,
                                      const int64_t* __restrict__ cap, int64_t cap_stride, int B, int V) {
    const int r = blockIdx.x, t = r / B, b = r - t * B;
    const int64_t tok = cap[(int64_t)b * cap_stride + t];
    if (tok < 0 || tok >= V) return;
    float* dst = table_grad + (size_t)tok * H;
    for (int h = threadIdx.x; h < H; h += blockDim.x) atomicAdd(dst + h, to_f32<TI>(dE[(int64_t)r * H + h]));
}

// set every row (b, t) of a (B, T, ld) T tensor's first D columns to src[b, :]
template <typename T>
__global__ void bcast_rows_kernel(T* __restrict__ dst, int64_t ld_t, int64_t ld_b, const T* __restrict__ src, int D, int Tn) {
    const int b = blockIdx.y, t = blockIdx.x;
    for (int d = threadIdx.x; d < D; d += blockDim.x) dst[(int64_t)b * ld_b + (int64_t)t * ld_t + d] = src[(size_t)b * D + d];
}

// LSTM cell backward pointwise (interleaved gates).  dh = dh_carry + dhd*mask.
template <typename T>
__global__ void lstm_bwd_kernel(const T* __restrict__ gates, int64_t ld_g, const float* __restrict__ c_prev, int64_t ld_cp,
                                const float* __restrict__ c_new, int64_t ld_cn, const float* __restrict__ dhd, int64_t ld_dhd,
                                const float* __restrict__ mask, int64_t ld_mask, const float* __restrict__ dh_carry,
                                float* __restrict__ dc_carry, T* __restrict__ dG, int64_t ld_dg, int B, int H) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x, b = blockIdx.y;
    pdl_wait();                       // dh_carry comes from the GEMM launched just before
    pdl_launch_dependents();
    if (j >= H) return;
    const T* g = gates + (int64_t)b * ld_g + 4 * j;
    const float gi = to_f32<T>(g[0]), gf = to_f32<T>(g[1]), gg = to_f32<T>(g[2]), go = to_f32<T>(g[3]);
    float dh = dh_carry[(size_t)b * H + j];
    if (dhd) dh += dhd[(int64_t)b * ld_dhd + j] * (mask ? mask[(int64_t)b * ld_mask + j] : 1.0f);
    const float cn = c_new[(int64_t)b * ld_cn + j];
    const float tc = tanhf(cn);
    const float d_o = dh * tc * go * (1.0f - go);
    const float dc = dh * go * (1.0f - tc * tc) + dc_carry[(size_t)b * H + j];
    const float d_i = dc * gg * gi * (1.0f - gi);
    const float d_f = dc * c_prev[(int64_t)b * ld_cp + j] * gf * (1.0f - gf);
    const float d_g = dc * gi * (1.0f - gg * gg);
    dc_carry[(size_t)b * H + j] = dc * gf;
    T* o = dG + (int64_t)b * ld_dg + 4 * j;
    o[0] = from_f32<T>(d_i); o[1] = from_f32<T>(d_f); o[2] = from_f32<T>(d_g); o[3] = from_f32<T>(d_o);
}

// column sums: out[n] = sum_m X[m, n]  (two-stage, deterministic)
template <typename T>
__global__ void colsum_partial_kernel(const T* __restrict__ X, int64_t ld, int M, int N, float* __restrict__ part, int rows_per_split) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    const int s = blockIdx.y;
    if (n >= N) return;
    const int m0 = s * rows_per_split, m1 = min(M, m0 + rows_per_split);
    float acc = 0.0f;
    for (int m = m0; m < m1; ++m) acc += to_f32<T>(X[(int64_t)m * ld + n]);
    part[(size_t)s * N + n] = acc;
}
// block = 8 row groups x 32 columns: 8 partial sums per column in flight, combined through shared memory (fixed order)
__global__ void colsum_final_kernel(const float* __restrict__ part, int S, int N, float* __restrict__ out, int ilH) {
    __shared__ float red[8][33];
    const int c = threadIdx.x & 31, rg = threadIdx.x >> 5;
    const int n = blockIdx.x * 32 + c;
    float acc = 0.0f;
    if (n < N)
        for (int s = rg; s < S; s += 8) acc += part[(size_t)s * N + n];
    red[rg][c] = acc;
    __syncthreads();
    if (rg == 0 && n < N) {
        float t = 0.0f;
#pragma unroll
        for (int r = 0; r < 8; ++r) t += red[r][c];
        // ilH > 0: column n of the interleaved gate axis belongs to source row (n%4)*ilH + n/4
        const int d = ilH > 0 ? ((n % 4) * ilH + n / 4) : n;
        out[d] = t;
    }
}
template <typename T>
int colsum(cudaStream_t s, const T* X, int64_t ld, int M, int N, float* out, float* scratch, int64_t scratch_floats, int ilH = 0) {
    int S = min(64, max(1, M / 64));
    while ((int64_t)S * N > scratch_floats && S > 1) S /= 2;
    SAT_REQUIRE((int64_t)S * N <= scratch_floats, "colsum scratch too small");
    const int rps = sat_cdiv(M, S);
    dim3 grid(sat_cdiv(N, 128), S);
    colsum_partial_kernel<T><<<grid, 128, 0, s>>>(X, ld, M, N, scratch, rps);
    colsum_final_kernel<<<sat_cdiv(N, 32), 256, 0, s>>>(scratch, S, N, out, ilH);
    SAT_CHECK_CUDA(cudaGetLastError());
    return SAT_OK;
}

// d(pre-tanh) of the initial state: dpre[b, j] = dh0 * (1 - h0^2), dpre[b, H + j] = dc0 * (1 - c0^2)
template <typename T>
__global__ void dinit_kernel(T* __restrict__ dpre, const float* __restrict__ dh0, const float* __restrict__ dc0,
                             const float* __restrict__ hc0, int B, int H) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B * H) return;
    const int b = i / H, j = i - b * H;
    const float h0 = hc0[(size_t)b * 2 * H + j], c0 = hc0[(size_t)b * 2 * H + H + j];
    dpre[(size_t)b * 2 * H + j] = from_f32<T>(dh0[i] * (1.0f - h0 * h0));
    dpre[(size_t)b * 2 * H + H + j] = from_f32<T>(dc0[i] * (1.0f - c0 * c0));
}

// split fp32 (B, 2H) [h0 | c0] into the state buffers
template <typename T>
__global__ void init_state_scatter_kernel(const float* __restrict__ hc0, T* __restrict__ h_dst, int64_t ld_h,
                                          float* __restrict__ c_dst, int64_t ld_c, int B, int H) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B * H) return;
    const int b = i / H, j = i - b * H;
    h_dst[(int64_t)b * ld_h + j] = from_f32<T>(hc0[(size_t)b * 2 * H + j]);
    c_dst[(int64_t)b * ld_c + j] = hc0[(size_t)b * 2 * H + H + j];
}

// argmax over V per row (first maximum), output int64 index   (decoder.py:132)
__global__ void argmax_rows_kernel(const float* __restrict__ x, int64_t ld, int V, int64_t* __restrict__ out, int64_t out_stride) {
    __shared__ float sv[32];
    __shared__ int si[32];
    const int r = blockIdx.x;
    const float* p = x + (int64_t)r * ld;
    float best = -INFINITY; int bi = 0x7fffffff;
    for (int v = threadIdx.x; v < V; v += blockDim.x) {
        const float f = p[v];
        if (f > best || (f == best && v < bi)) { best = f; bi = v; }
    }
    for (int o = 16; o > 0; o >>= 1) {
        const float ob = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ob > best || (ob == best && oi < bi)) { best = ob; bi = oi; }
    }
    if ((threadIdx.x & 31) == 0) { sv[threadIdx.x >> 5] = best; si[threadIdx.x >> 5] = bi; }
    __syncthreads();
    if (threadIdx.x < 32) {
        const int nw = blockDim.x >> 5;
        best = threadIdx.x < nw ? sv[threadIdx.x] : -INFINITY;
        bi = threadIdx.x < nw ? si[threadIdx.x] : 0x7fffffff;
        for (int o = 16; o > 0; o >>= 1) {
            const float ob = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ob > best || (ob == best && oi < bi)) { best = ob; bi = oi; }
        }
        if (threadIdx.x == 0) out[(int64_t)r * out_stride] = (bi == 0x7fffffff) ? 0 : bi;
    }
}

// ---- cross entropy over logits rows (train.py:150-151,161; Q10) ------------------------------------
// row r = b*T + t of logits (B*T, V).  Rows with t == T-1 are excluded.  rowloss[r] = lse - logit[target].
// Optionally writes dlogits = (softmax - onehot) * scale (T dtype), zero rows for t == T-1.  relu_mask != 0: the logits went
// through the final ReLU of the advanced deep output (decoder.py:158, Q8), so the gradient handed to the V-projection is
// zero wherever the logit is not positive.  A target outside [0, V) (the reference raises IndexError) sets *err_flag and
// contributes neither loss nor one-hot.
template <typename TO>
__global__ void ce_rows_kernel(const float* __restrict__ logits, int64_t ld, int V, int Tn, const int64_t* __restrict__ cap,
                               int64_t cap_stride, float* __restrict__ rowloss, TO* __restrict__ dlog, int64_t ld_d,
                               const float* __restrict__ scale_ptr, float scale_mul, int relu_mask, int* __restrict__ err_flag) {
    __shared__ float red[32];
    const int r = blockIdx.x, b = r / Tn, t = r - b * Tn;
    const int B = gridDim.x / Tn;
    const int64_t ro = (int64_t)t * B + b;      // dlogits are stored time-major
    const float* p = logits + (int64_t)r * ld;
    const bool active = (t < Tn - 1);
    if (!active) {
        if (rowloss && threadIdx.x == 0) rowloss[r] = 0.0f;
        if (dlog) for (int v = threadIdx.x; v < V; v += blockDim.x) dlog[ro * ld_d + v] = from_f32<TO>(0.0f);
        return;
    }
    float m = -INFINITY;
    for (int v = threadIdx.x; v < V; v += blockDim.x) m = fmaxf(m, p[v]);
    m = warp_max(m);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = m;
    __syncthreads();
    {
        float x = (threadIdx.x < (blockDim.x >> 5)) ? red[threadIdx.x] : -INFINITY;
        if (threadIdx.x < 32) { x = warp_max(x); if (threadIdx.x == 0) red[0] = x; }
    }
    __syncthreads();
    m = red[0];
    __syncthreads();
    float s = 0.0f;
    for (int v = threadIdx.x; v < V; v += blockDim.x) s += expf(p[v] - m);
    s = warp_sum(s);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    {
        float x = (threadIdx.x < (blockDim.x >> 5)) ? red[threadIdx.x] : 0.0f;
        if (threadIdx.x < 32) { x = warp_sum(x); if (threadIdx.x == 0) red[0] = x; }
    }
    __syncthreads();
    s = red[0];
    const int64_t tgt = cap[(int64_t)b * cap_stride + t + 1];
    const bool tgt_ok = (tgt >= 0 && tgt < V);
    if (!tgt_ok && err_flag && threadIdx.x == 0) atomicOr(err_flag, SAT_FLAG_BAD_TOKEN);
    const float lse = m + logf(s);
    if (rowloss && threadIdx.x == 0) rowloss[r] = tgt_ok ? lse - p[tgt] : 0.0f;
    if (dlog) {
        const float sc = (scale_ptr ? scale_ptr[0] : 1.0f) * scale_mul;
        const float inv = 1.0f / s;
        for (int v = threadIdx.x; v < V; v += blockDim.x) {
            const float x = p[v];
            float g = expf(x - m) * inv;
            if (v == tgt) g -= 1.0f;
            if (relu_mask && !(x > 0.0f)) g = 0.0f;
            dlog[ro * ld_d + v] = from_f32<TO>(g * sc);
        }
    }
}

// The same row pass for V <= 256 * CE_RPT with the row held in REGISTERS: one read of the logits (the three-pass kernel above reads
// a row three times and ran at 1.8 TB/s of useful bytes, 44 us at c2), two block reductions, one write of dlogits.
constexpr int CE_RPT = 16;
template <typename TO>
__global__ void __launch_bounds__(256) ce_rows_reg_kernel(const float* __restrict__ logits, int64_t ld, int V, int Tn,
                                                          const int64_t* __restrict__ cap, int64_t cap_stride, float* __restrict__ rowloss,
                                                          TO* __restrict__ dlog, int64_t ld_d, const float* __restrict__ scale_ptr,
                                                          float scale_mul, int relu_mask, int* __restrict__ err_flag) {
    __shared__ float red[8];
    __shared__ float bc;
    const int r = blockIdx.x, b = r / Tn, t = r - b * Tn;
    const int B = gridDim.x / Tn;
    const int64_t ro = (int64_t)t * B + b;      // dlogits are stored time-major
    const float* p = logits + (int64_t)r * ld;
    if (t >= Tn - 1) {                          // the last step is not part of the loss (Q10)
        if (rowloss && threadIdx.x == 0) rowloss[r] = 0.0f;
        if (dlog) for (int v = threadIdx.x; v < V; v += 256) dlog[ro * ld_d + v] = from_f32<TO>(0.0f);
        return;
    }
    float x[CE_RPT];
    float m = -INFINITY;
#pragma unroll
    for (int i = 0; i < CE_RPT; ++i) {
        const int v = threadIdx.x + 256 * i;
        x[i] = v < V ? __ldcs(p + v) : -INFINITY;
        m = fmaxf(m, x[i]);
    }
    m = warp_max(m);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x < 32) { float y = threadIdx.x < 8 ? red[threadIdx.x] : -INFINITY; y = warp_max(y); if (threadIdx.x == 0) bc = y; }
    __syncthreads();
    m = bc;
    float e[CE_RPT], s = 0.0f;
#pragma unroll
    for (int i = 0; i < CE_RPT; ++i) { e[i] = expf(x[i] - m); s += e[i]; }      // exp(-inf) = 0 for the padding
    s = warp_sum(s);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x < 32) { float y = threadIdx.x < 8 ? red[threadIdx.x] : 0.0f; y = warp_sum(y); if (threadIdx.x == 0) bc = y; }
    __syncthreads();
    s = bc;
    const int64_t tgt = cap[(int64_t)b * cap_stride + t + 1];
    const bool tgt_ok = (tgt >= 0 && tgt < V);
    if (!tgt_ok && err_flag && threadIdx.x == 0) atomicOr(err_flag, SAT_FLAG_BAD_TOKEN);
    if (rowloss && threadIdx.x == 0) rowloss[r] = tgt_ok ? m + logf(s) - p[tgt] : 0.0f;
    if (dlog) {
        const float sc = (scale_ptr ? scale_ptr[0] : 1.0f) * scale_mul;
        const float inv = 1.0f / s;
#pragma unroll
        for (int i = 0; i < CE_RPT; ++i) {
            const int v = threadIdx.x + 256 * i;
            if (v < V) {
                float g = e[i] * inv;
                if (v == tgt) g -= 1.0f;
                if (relu_mask && !(x[i] > 0.0f)) g = 0.0f;
                dlog[ro * ld_d + v] = from_f32<TO>(g * sc);
            }
        }
    }
}
template <typename TO>
inline void ce_rows(cudaStream_t st, int rows, const float* logits, int64_t ld, int V, int Tn, const int64_t* cap, int64_t cap_stride,
                    float* rowloss, TO* dlog, int64_t ld_d, const float* scale_ptr, float scale_mul, int relu_mask, int* err_flag) {
    if (V <= 256 * CE_RPT)
        ce_rows_reg_kernel<TO><<<rows, 256, 0, st>>>(logits, ld, V, Tn, cap, cap_stride, rowloss, dlog, ld_d, scale_ptr, scale_mul, relu_mask, err_flag);
    else
        ce_rows_kernel<TO><<<rows, 256, 0, st>>>(logits, ld, V, Tn, cap, cap_stride, rowloss, dlog, ld_d, scale_ptr, scale_mul, relu_mask, err_flag);
}

// loss[0] = ce mean, loss[1] = alpha_c * mean_{b,l} (1 - sum_t alpha)^2, loss[2] = total  (train.py:150-162).
// Two deterministic stages: LOSS_PARTS blocks each reduce a strided share of the CE rows and of the (b, l) columns into
// part[2 * block .. ], one block adds the partials up.
constexpr int LOSS_PARTS = 256;
__device__ __forceinline__ void block_sum2(float& x, float& y, float (*red)[32]) {
    x = warp_sum(x); y = warp_sum(y);
    if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = x; red[1][threadIdx.x >> 5] = y; }
    __syncthreads();
    const int nw = blockDim.x >> 5;
    x = (int)threadIdx.x < nw ? red[0][threadIdx.x] : 0.0f;
    y = (int)threadIdx.x < nw ? red[1][threadIdx.x] : 0.0f;
    if (threadIdx.x < 32) { x = warp_sum(x); y = warp_sum(y); }
}
__global__ void loss_partial_kernel(const float* __restrict__ rowloss, int rows, const float* __restrict__ alphas, int B, int Tn, int L,
                                    float* __restrict__ part) {
    __shared__ float red[2][32];
    float ce = 0.0f, reg = 0.0f;
    const int gtid = blockIdx.x * blockDim.x + threadIdx.x, gsz = gridDim.x * blockDim.x;
    for (int i = gtid; i < rows; i += gsz) ce += rowloss[i];
    if (alphas) {
        for (int i = gtid; i < B * L; i += gsz) {
            const int b = i / L, l = i - b * L;
            float s = 0.0f;
            for (int t = 0; t < Tn; ++t) s += alphas[((size_t)b * Tn + t) * L + l];
            reg += (1.0f - s) * (1.0f - s);
        }
    }
    block_sum2(ce, reg, red);
    if (threadIdx.x == 0) { part[2 * blockIdx.x] = ce; part[2 * blockIdx.x + 1] = reg; }
}
__global__ void loss_finalize_kernel(const float* __restrict__ part, int nparts, int n_active, int has_alphas, int B, int L,
                                     float alpha_c, float* __restrict__ loss, const int* __restrict__ err_flag) {
    __shared__ float red[2][32];
    float ce = 0.0f, reg = 0.0f;
    for (int i = threadIdx.x; i < nparts; i += blockDim.x) { ce += part[2 * i]; reg += part[2 * i + 1]; }
    block_sum2(ce, reg, red);
    if (threadIdx.x == 0) {
        loss[0] = n_active > 0 ? ce / (float)n_active : 0.0f;
        loss[1] = has_alphas ? alpha_c * reg / (float)(B * L) : 0.0f;
        loss[2] = loss[0] + loss[1];
        loss[3] = err_flag ? (float)err_flag[0] : 0.0f;      // SAT_FLAG_* bits of this call (0 = clean)
    }
}

// analytic grad of the doubly-stochastic regulariser: galpha[b,t,l] = 2 alpha_c (sum_t alpha - 1) / (B L) * scale  (Q11)
__global__ void alpha_reg_grad_kernel(const float* __restrict__ alphas, float* __restrict__ galpha, int B, int Tn, int L,
                                      float alpha_c, const float* __restrict__ scale_ptr) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B * L) return;
    const int b = i / L, l = i - b * L;
    float s = 0.0f;
    for (int t = 0; t < Tn; ++t) s += alphas[((size_t)b * Tn + t) * L + l];
    const float g = 2.0f * alpha_c * (s - 1.0f) / (float)(B * L) * (scale_ptr ? scale_ptr[0] : 1.0f);
    for (int t = 0; t < Tn; ++t) galpha[((size_t)b * Tn + t) * L + l] = g;
}

// out = relu'(ref) * x  (T)  ; used by the advanced-deep-output backward
template <typename T>
__global__ void relu_bwd_kernel(T* __restrict__ out, const float* __restrict__ x, const T* __restrict__ ref, int64_t n) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        out[i] = from_f32<T>(to_f32<T>(ref[i]) > 0.0f ? x[i] : 0.0f);
}
// dlog[t*B+b, v] (T, pitch ld_d) = dpreds[b,t,v] (* relu'(preds[b,t,v]) when ref != null: final ReLU of ado, Q8)
template <typename T>
__global__ void dpreds_to_dlog_kernel(T* __restrict__ out, int64_t ld_d, const float* __restrict__ x, const float* __restrict__ ref,
                                      int64_t ld_ref, int B, int Tn, int V) {
    const int r = blockIdx.x, b = r / Tn, t = r - b * Tn;
    const float* xr = x + (int64_t)r * V;
    const float* rr = ref ? ref + (int64_t)r * ld_ref : nullptr;
    T* o = out + ((int64_t)t * B + b) * ld_d;
    for (int v = threadIdx.x; v < V; v += blockDim.x) o[v] = from_f32<T>((rr && !(rr[v] > 0.0f)) ? 0.0f : xr[v]);
}
__global__ void fill_tokens_kernel(int64_t* __restrict__ p, int n, int64_t val) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = val;
}
// grad scatter with a flat token list: table_grad[tok[r], :] += dE[r, :]
template <typename TI>
__global__ void scatter_add_rows_kernel(float* __restrict__ table_grad, int H, const TI* __restrict__ dE,
                                        const int64_t* __restrict__ tok, int V) {
    const int r = blockIdx.x;
    const int64_t tk = tok[r];
    if (tk < 0 || tk >= V) return;
    float* dst = table_grad + (size_t)tk * H;
    for (int h = threadIdx.x; h < H; h += blockDim.x) atomicAdd(dst + h, to_f32<TI>(dE[(int64_t)r * H + h]));
}
__global__ void fill_kernel(float* __restrict__ p, int64_t n, float val) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) p[i] = val;
}
// y (T) = a (T) + b (T) + c (T)
template <typename T>
__global__ void add3_kernel(T* __restrict__ y, const T* __restrict__ a, const T* __restrict__ b, const T* __restrict__ c, int64_t n) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        y[i] = from_f32<T>(to_f32<T>(a[i]) + to_f32<T>(b[i]) + to_f32<T>(c[i]));
}
// y (T) += x (fp32)   elementwise into a T buffer
template <typename T>
__global__ void add_into_kernel(T* __restrict__ y, const float* __restrict__ x, int64_t n) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        y[i] = from_f32<T>(to_f32<T>(y[i]) + x[i]);
}
// loss[3] = status bits of a call that did not fuse the loss
__global__ void status_kernel(float* __restrict__ loss, const int* __restrict__ flags) {
    if (threadIdx.x == 0) { loss[0] = loss[1] = loss[2] = 0.0f; loss[3] = (float)flags[0]; }
}

// d img_features, the shares that are not a GEMM (see sat_decoder_backward).  One thread per (b, d) column keeps the T dctx
// values of its column in registers / L1 and walks the L locations.
//   mode 0: da[b,l,d] += dabar[b,d] / L                                   (generic attention kernels summed alpha (x) dctx)
//   mode 1: da[b,l,d] += sum_t alphas[b,t,l] dzh[t,b,d] + dabar[b,d] / L   (streaming attention kernels)
//   mode 2: da[b,l,d]  = (dabar[b,d] + sum_t (dzh[t,b,d] + extra[t,b,d])) / L   (no attention)
struct DfeatParams {
    float* da; const float* alphas; const float* dzh; int64_t ld_dzh; const float* dctx_extra; const float* dabar;
    int B, Tn, L, D, mode;
};
__global__ void __launch_bounds__(256) dfeat_kernel(const DfeatParams q) {
    const int d = blockIdx.x * blockDim.x + threadIdx.x, b = blockIdx.y;
    if (d >= q.D) return;
    const float invL = 1.0f / (float)q.L;
    const float mean_share = q.dabar[(size_t)b * q.D + d] * invL;
    float* col = q.da + (size_t)b * q.L * q.D + d;
    if (q.mode == 0) {
        for (int l = 0; l < q.L; ++l) col[(size_t)l * q.D] += mean_share;
    } else if (q.mode == 2) {
        float s = 0.0f;
        for (int t = 0; t < q.Tn; ++t) {
            s += q.dzh[((size_t)t * q.B + b) * q.ld_dzh + d];
            if (q.dctx_extra) s += q.dctx_extra[((size_t)t * q.B + b) * q.D + d];
        }
        const float v = mean_share + s * invL;
        for (int l = 0; l < q.L; ++l) col[(size_t)l * q.D] = v;
    } else {
        for (int l = 0; l < q.L; ++l) {
            float s = mean_share;
            for (int t = 0; t < q.Tn; ++t)       // alphas[b,t,l] is a warp-wide broadcast load
                s = fmaf(q.alphas[((size_t)b * q.Tn + t) * q.L + l], q.dzh[((size_t)t * q.B + b) * q.ld_dzh + d], s);
            col[(size_t)l * q.D] += s;
        }
    }
}

inline int grid_for(int64_t n) { return (int)(sat_cdiv64(n, 256) < (int64_t)(148 * 16) ? sat_cdiv64(n, 256) : (int64_t)(148 * 16)); }

}  // namespace misc
