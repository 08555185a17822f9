// Standalone tuning harness for the streaming attention kernels (not part of the library):
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo [-DSAT_ATTN_...=..] -o attn_bench tools/attn_bench.cu
//   ./attn_bench [B] [iters]
// Times fwd / bwd at D=2048, H=512, bf16 (config c2) with CUDA events; inputs alternate between two sets (> L2).
#include <stdarg.h>
#include <stdlib.h>
#include <vector>
#include "../show-attend-and-tell_b200/csrc/sat_attn_stream.cuh"

int g_attn_generic = 0;
int g_attn_part = 0;
long long g_attn_pin_bytes = 0;
int g_attn_wa_hint_fwd = 1, g_attn_wa_hint_bwd = 0;
int g_pdl = 1;
void sat_set_error(const char* fmt, ...) { va_list ap; va_start(ap, fmt); vfprintf(stderr, fmt, ap); va_end(ap); fprintf(stderr, "\n"); }
SatProfScope::SatProfScope(int c, cudaStream_t s) : cat(c), st(s), slot(-1) {}
SatProfScope::~SatProfScope() {}

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { fprintf(stderr, "%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e)); exit(1); } } while (0)

__global__ void fill_bf16(bf16* p, size_t n, uint32_t seed, float scale, int relu) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        uint32_t x = (uint32_t)i * 2654435761u ^ seed; x ^= x >> 15; x *= 2246822519u; x ^= x >> 13;
        float f = ((x & 0xffff) / 32768.0f - 1.0f) * scale;
        if (relu && f < 0) f = 0;
        p[i] = __float2bfloat16(f);
    }
}
__global__ void fill_f32(float* p, size_t n, uint32_t seed, float scale) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        uint32_t x = (uint32_t)i * 2654435761u ^ seed; x ^= x >> 15; x *= 2246822519u; x ^= x >> 13;
        p[i] = ((x & 0xffff) / 32768.0f - 1.0f) * scale;
    }
}

__global__ void read_only(const uint4* __restrict__ x, size_t n, uint32_t* out) {
    uint32_t acc = 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const uint4 t = __ldcs(x + i); acc ^= t.x ^ t.y ^ t.z ^ t.w;
    }
    if (acc == 0x12345678u) *out = acc;
}
__global__ void checksum_f32(const float* x, size_t n, double* out) {
    double acc = 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) acc += (double)x[i] * (1.0 + (i % 7));
    atomicAdd(out, acc);
}
__global__ void checksum_bf16(const bf16* x, size_t n, double* out) {
    double acc = 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) acc += (double)__bfloat162float(x[i]) * (1.0 + (i % 7));
    atomicAdd(out, acc);
}

int main(int argc, char** argv) {
    if (getenv("SAT_PART")) g_attn_part = atoi(getenv("SAT_PART"));
    if (getenv("SAT_PIN_MB")) g_attn_pin_bytes = (long long)atoi(getenv("SAT_PIN_MB")) << 20;
    if (getenv("SAT_PDL")) g_pdl = atoi(getenv("SAT_PDL"));
    if (getenv("SAT_WA_FWD")) g_attn_wa_hint_fwd = atoi(getenv("SAT_WA_FWD"));
    if (getenv("SAT_WA_BWD")) g_attn_wa_hint_bwd = atoi(getenv("SAT_WA_BWD"));
    const int static_in = getenv("SAT_STATIC") ? atoi(getenv("SAT_STATIC")) : 1;
    const int sameset = getenv("SAT_SAMESET") ? atoi(getenv("SAT_SAMESET")) : 0;      // 1: every launch reads the same inputs
    const int B = argc > 1 ? atoi(argv[1]) : 256, iters = argc > 2 ? atoi(argv[2]) : 20;
    const int L = 196, D = 2048, H = 512;
    bf16 *a[2], *wa[2], *ctx, *z, *duf, *dpre;
    float *uf, *v, *bv, *alpha, *ctx32, *dz, *galpha, *dv, *dbv;
    for (int s = 0; s < 2; ++s) {
        CK(cudaMalloc(&a[s], (size_t)B * L * D * 2)); CK(cudaMalloc(&wa[s], (size_t)B * L * H * 2));
        fill_bf16<<<1024, 256>>>(a[s], (size_t)B * L * D, 17 + s, 1.0f, 1);
        fill_bf16<<<1024, 256>>>(wa[s], (size_t)B * L * H, 91 + s, 1.0f, 0);
    }
    CK(cudaMalloc(&ctx, (size_t)B * D * 2)); CK(cudaMalloc(&z, (size_t)B * D * 2)); CK(cudaMalloc(&duf, (size_t)B * (H + D) * 2));
    CK(cudaMalloc(&dpre, (size_t)B * L * H * 2));
    CK(cudaMalloc(&uf, (size_t)B * (H + D) * 4)); CK(cudaMalloc(&v, H * 4)); CK(cudaMalloc(&bv, 4)); CK(cudaMalloc(&alpha, (size_t)B * L * 4));
    CK(cudaMalloc(&ctx32, (size_t)B * D * 4)); CK(cudaMalloc(&dz, (size_t)B * D * 4)); CK(cudaMalloc(&galpha, (size_t)B * L * 4));
    CK(cudaMalloc(&dv, (size_t)B * H * 4)); CK(cudaMalloc(&dbv, (size_t)B * 4));
    fill_f32<<<256, 256>>>(uf, (size_t)B * (H + D), 5, 1.0f); fill_f32<<<4, 256>>>(v, H, 6, 0.05f); CK(cudaMemset(bv, 0, 4));
    fill_f32<<<256, 256>>>(dz, (size_t)B * D, 7, 1.0f); fill_f32<<<64, 256>>>(galpha, (size_t)B * L, 8, 0.01f);
    const int64_t sb = attn_s::scratch_bytes(B, 1, D, H);
    uint8_t* scratch; CK(cudaMalloc(&scratch, sb)); CK(cudaMemset(scratch, 0, sb));
    CK(cudaDeviceSynchronize());

    auto fwd = [&](int s) {
        attn::FwdParams p; memset(&p, 0, sizeof(p));
        p.a = a[s]; p.wa = wa[s]; p.uf = uf; p.ld_uf = H + D; p.v = v; p.bv = bv; p.alpha = alpha; p.ld_alpha = L;
        p.ctx = ctx; p.ld_ctx = D; p.z = z; p.ld_z = D; p.NI = B; p.NB = 1; p.L = L; p.D = D; p.H = H; p.gate = 1;
        p.ctx32 = ctx32; p.ld_c32 = D; p.s_part = (float*)scratch; p.static_inputs = static_in;
        if (attn_s::launch_fwd<bf16, false>(p, 0) != 0) exit(2);
    };
    auto bwd = [&](int s) {
        attn::BwdParams p; memset(&p, 0, sizeof(p));
        p.a = a[s]; p.wa = wa[s]; p.uf = uf; p.ld_uf = H + D; p.v = v; p.alpha = alpha; p.ld_alpha = L; p.ctx = ctx; p.ld_ctx = D;
        p.dz = dz; p.ld_dz = D; p.galpha = galpha; p.ld_ga = L; p.duf = duf; p.ld_duf = H + D; p.dv_acc = dv; p.dbv_acc = dbv;
        p.first = 1; p.B = B; p.L = L; p.D = D; p.H = H; p.gate = 1; p.dpre_out = dpre; p.ctx32 = ctx32; p.ld_c32 = D;
        p.s_part = (float*)scratch; p.dz_is_dctx = 1; p.static_inputs = static_in;
        if (attn_s::launch_bwd<bf16, false>(p, 0) != 0) exit(2);
    };
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    setvbuf(stdout, nullptr, _IONBF, 0);
    {   // calibration of the box: device-to-device copy of 512 MB (read + write bytes), best of 5
        uint8_t *x, *y; const size_t n = (size_t)512 << 20;
        CK(cudaMalloc(&x, n)); CK(cudaMalloc(&y, n)); CK(cudaMemset(x, 1, n));
        float best = 1e9f;
        for (int i = 0; i < 6; ++i) {
            CK(cudaEventRecord(e0)); CK(cudaMemcpyAsync(y, x, n, cudaMemcpyDeviceToDevice)); CK(cudaEventRecord(e1));
            CK(cudaEventSynchronize(e1)); float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); if (i && ms < best) best = ms;
        }
        printf("calib copy %.0f GB/s  ", 2.0 * n / best / 1e6);
        uint32_t* o; CK(cudaMalloc(&o, 4)); best = 1e9f;
        for (int i = 0; i < 6; ++i) {
            CK(cudaEventRecord(e0)); read_only<<<148 * 8, 512>>>((const uint4*)x, n / 16, o); CK(cudaEventRecord(e1));
            CK(cudaEventSynchronize(e1)); float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); if (i && ms < best) best = ms;
        }
        printf("read-only %.0f GB/s  ", 1.0 * n / best / 1e6);
        CK(cudaFree(x)); CK(cudaFree(y));
    }
    const int only = argc > 3 ? atoi(argv[3]) : -1;
    for (int which = 0; which < 2; ++which) {
        if (only >= 0 && which != only) continue;
        for (int i = 0; i < 3; ++i) { if (which) bwd(sameset ? 0 : i & 1); else fwd(sameset ? 0 : i & 1); }
        CK(cudaDeviceSynchronize());
        CK(cudaEventRecord(e0));
        for (int i = 0; i < iters; ++i) { if (which) bwd(sameset ? 0 : i & 1); else fwd(sameset ? 0 : i & 1); }
        CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        const double us = ms * 1e3 / iters;
        const double bytes = which ? (double)B * (L * (H + D) * 2.0 + 2.0 * L * H * 4) : (double)B * L * (H + D) * 2.0;
        const double real = which ? (double)B * (L * (H + D) * 2.0 + L * H * 2.0) : bytes;
        printf("%s B=%d  %.2f us  algorithmic %.0f GB/s  moved %.0f GB/s\n", which ? "bwd" : "fwd", B, us, bytes / us / 1e3, real / us / 1e3);
    }
    CK(cudaDeviceSynchronize());
    if (getenv("SAT_TRAIN")) {   // the access pattern of a train step: T forward launches then T backward launches on one batch, next batch ...
        const int T = 20;
        cudaEvent_t ef0, ef1, eb1; CK(cudaEventCreate(&ef0)); CK(cudaEventCreate(&ef1)); CK(cudaEventCreate(&eb1));
        double tf = 0, tb = 0;
        for (int it = 0; it < 6; ++it) {
            const int s = it & 1;
            CK(cudaEventRecord(ef0));
            for (int t = 0; t < T; ++t) fwd(s);
            CK(cudaEventRecord(ef1));
            for (int t = 0; t < T; ++t) bwd(s);
            CK(cudaEventRecord(eb1)); CK(cudaEventSynchronize(eb1));
            float a, b; CK(cudaEventElapsedTime(&a, ef0, ef1)); CK(cudaEventElapsedTime(&b, ef1, eb1));
            if (it >= 2) { tf += a; tb += b; }
        }
        printf("train pattern B=%d wa_hint fwd=%d bwd=%d: fwd %.2f us/launch  bwd %.2f us/launch\n", B, g_attn_wa_hint_fwd, g_attn_wa_hint_bwd,
               tf * 1e3 / (4 * T), tb * 1e3 / (4 * T));
    }
    {   // checksums of the last fwd / bwd outputs: compare across SAT_PART settings
        double* cs; CK(cudaMalloc(&cs, 5 * 8)); CK(cudaMemset(cs, 0, 5 * 8));
        fwd(0); bwd(0);
        checksum_f32<<<64, 256>>>(ctx32, (size_t)B * D, cs); checksum_f32<<<64, 256>>>(alpha, (size_t)B * L, cs + 1);
        checksum_bf16<<<64, 256>>>(z, (size_t)B * D, cs + 2); checksum_f32<<<64, 256>>>(dv, (size_t)B * H, cs + 3);
        checksum_f32<<<8, 256>>>(dbv, (size_t)B, cs + 4);
        double h[5]; CK(cudaMemcpy(h, cs, sizeof(h), cudaMemcpyDeviceToHost));
        printf("checksums part=%d: ctx32 %.6f alpha %.6f z %.4f dv %.6f dbv %.6f\n", g_attn_part, h[0], h[1], h[2], h[3], h[4]);
    }
    return 0;
}
