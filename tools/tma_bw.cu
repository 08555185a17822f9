// Microbenchmark (not part of the library): how fast can ONE SM pull a contiguous stream out of L2/HBM?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tma_bw tools/tma_bw.cu
//   ./tma_bw G slot_bytes nslots producers chunk_bytes [MB_per_cta]
// ring mode: `producers` elected threads (one per warp) issue cp.async.bulk into a ring of `nslots` slots (slot n is
// issued by producer n % producers, split into chunk_bytes requests); one consumer thread per producer waits for the
// slot and hands it straight back.  ldg mode (producers == 0): 256 threads read with 16-byte loads, 8 in flight each.
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>
#include "../show-attend-and-tell_b200/csrc/sat_common.cuh"

void sat_set_error(const char*, ...) {}
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { fprintf(stderr, "%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e)); exit(1); } } while (0)

__global__ void __launch_bounds__(512, 1)
ring_kernel(const uint8_t* __restrict__ src, size_t per_cta, int slot_bytes, int nslots, int P, int chunk) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t* full = reinterpret_cast<uint64_t*>(smem);
    uint64_t* empty = full + 64;
    uint8_t* ring = smem + 1024;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) {
        for (int i = 0; i < nslots; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
        mbar_fence_init();
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();
    const uint8_t* base = src + (size_t)blockIdx.x * per_cta;
    const int total = (int)(per_cta / slot_bytes);
    if (lane != 0) return;
    if (warp < P) {                 // producer `warp`
        for (int n = warp; n < total; n += P) {
            const int r = n % nslots, ph = (n / nslots) & 1;
            mbar_wait(&empty[r], ph ^ 1);
            mbar_expect_tx(&full[r], (uint32_t)slot_bytes);
            for (int o = 0; o < slot_bytes; o += chunk)
                bulk_g2s(ring + (size_t)r * slot_bytes + o, base + (size_t)n * slot_bytes + o, (uint32_t)min(chunk, slot_bytes - o), &full[r]);
        }
    } else if (warp < 2 * P) {      // consumer of the same slots
        const int w = warp - P;
        for (int n = w; n < total; n += P) {
            const int r = n % nslots, ph = (n / nslots) & 1;
            mbar_wait(&full[r], ph);
            mbar_arrive(&empty[r]);
        }
    }
}

__global__ void __launch_bounds__(256, 1)
ldg_kernel(const uint4* __restrict__ src, size_t per_cta16, uint32_t* out) {
    const uint4* base = src + (size_t)blockIdx.x * per_cta16;
    uint32_t acc = 0;
    for (size_t i = threadIdx.x; i + 7 * 256 < per_cta16; i += 8 * 256) {
        uint4 t[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) t[k] = __ldcs(base + i + k * 256);
#pragma unroll
        for (int k = 0; k < 8; ++k) acc ^= t[k].x ^ t[k].y ^ t[k].z ^ t[k].w;
    }
    if (acc == 0x12345678u) *out = acc;
}

int main(int argc, char** argv) {
    const int G = argc > 1 ? atoi(argv[1]) : 148, slot = argc > 2 ? atoi(argv[2]) : 10240, nslots = argc > 3 ? atoi(argv[3]) : 17;
    const int P = argc > 4 ? atoi(argv[4]) : 1, chunk = argc > 5 ? atoi(argv[5]) : slot;
    const size_t mb = argc > 6 ? atoi(argv[6]) : 4;
    const size_t per_cta = (mb << 20) / slot * slot;
    uint8_t* src; CK(cudaMalloc(&src, per_cta * G + 4096)); CK(cudaMemset(src, 1, per_cta * G));
    uint32_t* out; CK(cudaMalloc(&out, 4));
    const size_t smem = 1024 + (size_t)slot * nslots;
    CK(cudaFuncSetAttribute(ring_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    float best = 1e9f;
    for (int it = 0; it < 5; ++it) {
        CK(cudaEventRecord(e0));
        if (P > 0) ring_kernel<<<G, 32 * 2 * P, smem>>>(src, per_cta, slot, nslots, P, chunk);
        else ldg_kernel<<<G, 256>>>((const uint4*)src, per_cta / 16, out);
        CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); CK(cudaGetLastError());
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); if (it && ms < best) best = ms;
    }
    printf("G=%d slot=%d nslots=%d P=%d chunk=%d per_cta=%zuMB: %.1f us  %.1f GB/s per SM  %.0f GB/s total\n", G, slot, nslots, P, chunk, mb,
           best * 1e3, per_cta / best / 1e6, per_cta * G / best / 1e6);
    return 0;
}
