// Standalone tuning harness for the tcgen05 GEMM on the decoder's per-step (skinny, M = batch) shapes -- not part of the
// library.   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -lcuda -o gemm_bench tools/gemm_bench.cu
//   ./gemm_bench [iters]
#include <stdarg.h>
#include <stdlib.h>
#include <vector>
#include "../show-attend-and-tell_b200/csrc/sat_gemm_tc.cuh"
#include "../show-attend-and-tell_b200/csrc/sat_epilogue.cuh"

int g_pdl = 1;
int g_gemm_max_split = 4;
void sat_set_error(const char* fmt, ...) { va_list ap; va_start(ap, fmt); vfprintf(stderr, fmt, ap); va_end(ap); fprintf(stderr, "\n"); }
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { fprintf(stderr, "%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e)); exit(1); } } while (0)

__global__ void fill_bf16(bf16* p, size_t n, uint32_t seed) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        uint32_t x = (uint32_t)i * 2654435761u ^ seed; x ^= x >> 15; x *= 2246822519u; x ^= x >> 13;
        p[i] = __float2bfloat16(((x & 0xffff) / 32768.0f - 1.0f) * 0.25f);
    }
}
__global__ void checksum_f32(const float* x, size_t n, double* out) {
    double acc = 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) acc += (double)x[i] * (1.0 + (i % 7));
    atomicAdd(out, acc);
}

struct Shape { const char* name; int M, N, K; bool b_kmajor; bool out_bf16; int ldc; bool a_mn; };

int main(int argc, char** argv) {
    const int iters = argc > 1 ? atoi(argv[1]) : 50;
    if (getenv("SAT_PDL")) g_pdl = atoi(getenv("SAT_PDL"));
    if (getenv("SAT_MAXSPLIT")) g_gemm_max_split = atoi(getenv("SAT_MAXSPLIT"));
    const Shape shapes[] = {
        {"UF   = h [U;f_beta]^T        ", 256, 2560, 512, true},
        {"gates= [z|h] W_gzh^T         ", 256, 2048, 2560, true},
        {"dZH  = dG W_gzh     (B MN)   ", 256, 2560, 2048, false},
        {"dh   = dUF w_uf     (B MN)   ", 256, 512, 2560, false},
        {"c1 UF (B=32)                 ", 32, 1024, 512, true},
        {"Wa   = a W^T (K1), bf16 out  ", 50176, 512, 2048, true, true},
        {"dW_att = dWa^T a (A,B MN)    ", 512, 2048, 50176, false, false, 0, true},
        {"dW_gzh = dG^T ZH (A,B MN)    ", 2048, 2560, 5120, false, false, 0, true},
        {"logits = Hdrop W_out^T, ld=V ", 5120, 2531, 512, true},
        {"logits, ld=Vp (train path)   ", 5120, 2531, 512, true, false, 2536},
        {"probe: one tile, one k-block ", 128, 32, 64, true},
        {"probe: 128 tiles, 1 k-block  ", 256, 2048, 64, true},
        {"probe: 128 tiles, 8 k-blocks ", 256, 2048, 512, true},
        {"probe: 128 tiles, 16 k-blocks", 256, 2048, 1024, true},
        {"probe: 128 tiles, 80 k-blocks", 256, 2048, 5120, true},
        {"probe: 16 tiles, 8 k-blocks  ", 256, 256, 512, true},
    };
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    double* cs; CK(cudaMalloc(&cs, 8));
    for (const Shape& s : shapes) {
        bf16 *A, *B; float* C;
        const int Np = (s.N + 7) & ~7;
        CK(cudaMalloc(&A, (size_t)s.M * s.K * 2)); CK(cudaMalloc(&B, (size_t)Np * s.K * 2)); const int ldc = s.ldc ? s.ldc : s.N;
        CK(cudaMalloc(&C, (size_t)s.M * ldc * 4));
        CK(cudaMemset(C, 0, (size_t)s.M * ldc * 4));
        fill_bf16<<<512, 256>>>(A, (size_t)s.M * s.K, 3); fill_bf16<<<512, 256>>>(B, (size_t)Np * s.K, 5);
        EpiLinear<bf16, false> e;
        memset(&e, 0, sizeof(e));
        if (s.out_bf16) { e.Ct = reinterpret_cast<bf16*>(C); e.ldct = ldc; } else { e.Cf = C; e.ldcf = ldc; }
        e.M = s.M; e.N = s.N; e.vec_ok = (ldc % 8 == 0);
        const int64_t ldb = s.b_kmajor ? s.K : s.N;
        const int64_t lda = s.a_mn ? s.M : s.K;
        const int bn_over = getenv("SAT_BN") ? atoi(getenv("SAT_BN")) : 0;
        const int sp_over = getenv("SAT_SPLIT") ? atoi(getenv("SAT_SPLIT")) : 0;
        auto run = [&]() { if (tc::launch(0, A, lda, !s.a_mn, B, ldb, s.b_kmajor, s.M, s.N, s.K, e, bn_over, sp_over) != 0) exit(2); };
        for (int i = 0; i < 3; ++i) run();
        CK(cudaDeviceSynchronize());
        // (a) eager launches: includes the host cost of two tensor-map encodes + launch per call
        CK(cudaEventRecord(e0));
        for (int i = 0; i < iters; ++i) run();
        CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        float ms_eager; CK(cudaEventElapsedTime(&ms_eager, e0, e1));
        // (b) the same launches replayed from a CUDA graph: device time only
        cudaStream_t st; CK(cudaStreamCreate(&st));
        cudaGraph_t g; cudaGraphExec_t ge;
        CK(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
        for (int i = 0; i < iters; ++i) if (tc::launch(st, A, lda, !s.a_mn, B, ldb, s.b_kmajor, s.M, s.N, s.K, e, bn_over, sp_over) != 0) exit(2);
        CK(cudaStreamEndCapture(st, &g));
        CK(cudaGraphInstantiate(&ge, g, 0));
        CK(cudaGraphLaunch(ge, st)); CK(cudaStreamSynchronize(st));
        CK(cudaEventRecord(e0, st)); CK(cudaGraphLaunch(ge, st)); CK(cudaEventRecord(e1, st)); CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        CK(cudaGraphExecDestroy(ge)); CK(cudaGraphDestroy(g)); CK(cudaStreamDestroy(st));
        CK(cudaMemset(cs, 0, 8));
        checksum_f32<<<64, 256>>>(C, (size_t)s.M * ldc / (s.out_bf16 ? 2 : 1), cs);
        double h; CK(cudaMemcpy(&h, cs, 8, cudaMemcpyDeviceToHost));
        const double us = ms * 1e3 / iters;
        printf("%s M=%5d N=%4d K=%4d  graph %7.2f us  eager %7.2f us  %7.1f TFLOP/s  checksum %.6e\n", s.name, s.M, s.N, s.K, us,
               ms_eager * 1e3 / iters, 2.0 * s.M * s.N * s.K / us / 1e6, h);
        CK(cudaFree(A)); CK(cudaFree(B)); CK(cudaFree(C));
    }
    return 0;
}
