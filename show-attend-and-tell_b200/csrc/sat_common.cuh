// Common device/host helpers for the sat_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

typedef __nv_bfloat16 bf16;

// ---------------------------------------------------------------------------------------------
// error plumbing (thread-local last-error string, returned through sat_last_error())
// ---------------------------------------------------------------------------------------------
void sat_set_error(const char* fmt, ...);
#define SAT_OK 0
#define SAT_ERR_ARG -1
#define SAT_ERR_CUDA -2
#define SAT_ERR_UNSUPPORTED -3
// device-side status bits of a sequence-driver call (sat_decoder_args.loss[3], sat_decoder_status)
#define SAT_FLAG_BAD_TOKEN 1     /* a caption id outside [0, V): the reference raises IndexError (nn.Embedding / CrossEntropyLoss) */

#define SAT_CHECK_CUDA(expr)                                                                  \
    do {                                                                                      \
        cudaError_t _e = (expr);                                                              \
        if (_e != cudaSuccess) {                                                              \
            sat_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
            return SAT_ERR_CUDA;                                                              \
        }                                                                                     \
    } while (0)

#define SAT_REQUIRE(cond, ...)                                                                \
    do {                                                                                      \
        if (!(cond)) {                                                                        \
            sat_set_error(__VA_ARGS__);                                                       \
            return SAT_ERR_ARG;                                                               \
        }                                                                                     \
    } while (0)

#define SAT_TRY(expr)                                                                         \
    do {                                                                                      \
        int _r = (expr);                                                                      \
        if (_r != SAT_OK) return _r;                                                          \
    } while (0)

// Kernel attributes (cudaFuncSetAttribute) belong to a device context, not to the process: a function-local `static bool`
// would configure only the first device a process touches.  One of these per call site; `first()` is true exactly once per
// device (thread-safe: the caller configures while holding the lock).
#include <mutex>
struct SatPerDevice {
    std::mutex mu;
    unsigned long long done = 0;
    template <typename F> int once(F&& fn) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess) dev = 0;
        const unsigned long long bit = 1ull << (dev & 63);
        std::lock_guard<std::mutex> lk(mu);
        if (done & bit) return SAT_OK;
        const int r = fn();
        if (r == SAT_OK) done |= bit;
        return r;
    }
};

static inline int sat_cdiv(int a, int b) { return (a + b - 1) / b; }
static inline int64_t sat_cdiv64(int64_t a, int64_t b) { return (a + b - 1) / b; }

// ---------------------------------------------------------------------------------------------
// scalar conversion helpers
// ---------------------------------------------------------------------------------------------
template <typename T> __device__ __forceinline__ float to_f32(T x);
template <> __device__ __forceinline__ float to_f32<float>(float x) { return x; }
template <> __device__ __forceinline__ float to_f32<bf16>(bf16 x) { return __bfloat162float(x); }
template <typename T> __device__ __forceinline__ T from_f32(float x);
template <> __device__ __forceinline__ float from_f32<float>(float x) { return x; }
template <> __device__ __forceinline__ bf16 from_f32<bf16>(float x) { return __float2bfloat16_rn(x); }

// 16-byte vector of T (4 floats or 8 bf16) unpacked to fp32
template <typename T> struct Vec16;
template <> struct Vec16<float> {
    static constexpr int N = 4;
    __device__ __forceinline__ static void unpack(const uint4& u, float* f) {
        f[0] = __uint_as_float(u.x); f[1] = __uint_as_float(u.y);
        f[2] = __uint_as_float(u.z); f[3] = __uint_as_float(u.w);
    }
    __device__ __forceinline__ static uint4 pack(const float* f) {
        return make_uint4(__float_as_uint(f[0]), __float_as_uint(f[1]), __float_as_uint(f[2]), __float_as_uint(f[3]));
    }
};
template <> struct Vec16<bf16> {
    static constexpr int N = 8;
    __device__ __forceinline__ static void unpack(const uint4& u, float* f) {
        const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            f[2 * i] = __uint_as_float(w[i] << 16);
            f[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
        }
    }
    __device__ __forceinline__ static uint4 pack(const float* f) {
        uint32_t w[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            __nv_bfloat162 h = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
            w[i] = *reinterpret_cast<uint32_t*>(&h);
        }
        return make_uint4(w[0], w[1], w[2], w[3]);
    }
};

// 4 consecutive T elements <-> fp32 (8-byte access for bf16, 16-byte for fp32; pointer must be aligned accordingly)
template <typename T> __device__ __forceinline__ void store4(T* p, const float* f);
template <> __device__ __forceinline__ void store4<float>(float* p, const float* f) {
    *reinterpret_cast<float4*>(p) = make_float4(f[0], f[1], f[2], f[3]);
}
template <> __device__ __forceinline__ void store4<bf16>(bf16* p, const float* f) {
    __nv_bfloat162 lo = __floats2bfloat162_rn(f[0], f[1]), hi = __floats2bfloat162_rn(f[2], f[3]);
    *reinterpret_cast<uint2*>(p) = make_uint2(*reinterpret_cast<uint32_t*>(&lo), *reinterpret_cast<uint32_t*>(&hi));
}
template <typename T> __device__ __forceinline__ void load4(const T* p, float* f);
template <> __device__ __forceinline__ void load4<float>(const float* p, float* f) {
    const float4 v = *reinterpret_cast<const float4*>(p);
    f[0] = v.x; f[1] = v.y; f[2] = v.z; f[3] = v.w;
}
template <> __device__ __forceinline__ void load4<bf16>(const bf16* p, float* f) {
    const uint2 u = *reinterpret_cast<const uint2*>(p);
    f[0] = __uint_as_float(u.x << 16); f[1] = __uint_as_float(u.x & 0xffff0000u);
    f[2] = __uint_as_float(u.y << 16); f[3] = __uint_as_float(u.y & 0xffff0000u);
}

__device__ __forceinline__ uint4 ldg_stream16(const void* p) {   // streaming 128-bit load, no L1 allocate
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}

// ---------------------------------------------------------------------------------------------
// math
// ---------------------------------------------------------------------------------------------
template <bool PRECISE> __device__ __forceinline__ float sat_tanh(float x) {
    if (PRECISE) return tanhf(x);
    float y;
    asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
// tanh of the attention scores e_l = v . tanh(Wa_l + Uh) (attention.py:17-18).  fp32 mode: the score step of a 128-image beam-5
// search evaluates 64 M tanh per decode step, and libm's tanhf (~25 instructions, two branches) made the fp32 attention kernel
// compute-bound (93 us per step).  1 - 2 / (1 + e^{2x}) with ex2.approx + rcp.approx is 2 MUFU + 4 FMA-pipe instructions; its
// ABSOLUTE error is <= ~3e-7 (saturates correctly to +-1), which is what matters here: the tanh feeds a dot product with |v| < 0.05,
// so a score moves by ~1e-7 -- the size of libm-vs-libm differences -- while relative accuracy near 0 is irrelevant.  The LSTM cell
// and the initial state keep tanhf.
template <bool PRECISE> __device__ __forceinline__ float attn_tanh(float x) {
    if (!PRECISE) return sat_tanh<false>(x);
    float e, r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(x * 2.885390081777927f));      // e^{2x} = 2^{2x log2 e}
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(1.0f + e));
    return fmaf(-2.0f, r, 1.0f);
}
// tanh(w + u) from e^{2w} and e^{2u} kept separately: when one w meets several u (the beam rows of an image) the exponential of
// w is taken once and each pair costs one reciprocal -- 1 + 1/NB SFU operations per element instead of 2.  The exponents are
// clamped to +-40 so that neither factor is 0 or inf (their product may be: tanh then saturates correctly, never NaN); the clamp
// only matters for |w| > 40 with u of the opposite sign within ~10 of it, far outside what a tanh pre-activation sees.
__device__ __forceinline__ float exp2x_clamped(float x) {
    float e;
    const float a = fminf(fmaxf(x * 2.885390081777927f, -115.4f), 115.4f);
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(a));
    return e;
}
__device__ __forceinline__ float tanh_from_exp2x(float e) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(1.0f + e));
    return fmaf(-2.0f, r, 1.0f);
}
template <bool PRECISE> __device__ __forceinline__ float sat_sigmoid(float x) {
    if (PRECISE) return 1.0f / (1.0f + expf(-x));
    return 0.5f * sat_tanh<false>(0.5f * x) + 0.5f;
}
template <bool PRECISE> __device__ __forceinline__ float sat_exp(float x) {
    return PRECISE ? expf(x) : __expf(x);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// ---------------------------------------------------------------------------------------------
// PTX wrappers: shared addresses, mbarrier, bulk copies (TMA), clusters, tcgen05
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) {}
}

// 1-D bulk async copy global -> shared (TMA engine, UBLKCP), completion on an mbarrier.
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
// same with an L2 eviction-priority hint (createpolicy encodings: evict_first for data read once, evict_last for data
// worth keeping in the 126 MB L2 until its next use)
constexpr uint64_t L2_EVICT_NORMAL = 0x1000000000000000ull, L2_EVICT_FIRST = 0x12F0000000000000ull, L2_EVICT_LAST = 0x14F0000000000000ull;
__device__ __forceinline__ void bulk_g2s_hint(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar, uint64_t policy) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                 ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)), "l"(policy) : "memory");
}
// 2-D tiled TMA load (UTMALDG)
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(smem_dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1) : "memory");
}
// same with an L2 eviction-priority hint
__device__ __forceinline__ void tma_load_2d_hint(void* smem_dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar, uint64_t policy) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%3, %4}], [%2], %5;"
        ::"r"(smem_u32(smem_dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "l"(policy) : "memory");
}
// pull one 128-byte line into L2 without waiting for it
__device__ __forceinline__ void l2_prefetch(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
// ask L2 for `bytes` (multiple of 16) starting at p, asynchronously, without a destination (TMA engine, no shared memory)
__device__ __forceinline__ void l2_prefetch_bulk(const void* p, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}

// Programmatic dependent launch (PDL).  A kernel launched through pdl_config() may start while its predecessor in the
// stream is still running; it must execute pdl_wait() before its first access to global memory the predecessor (or any
// earlier kernel) wrote or still reads.  Both instructions are no-ops in a kernel launched the ordinary way.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
extern int g_pdl;              // sat_set_pdl: 1 (default) = launch PDL-aware kernels with programmatic stream serialization
inline void pdl_config(cudaLaunchConfig_t* cfg, cudaLaunchAttribute* attr, dim3 grid, dim3 block, size_t smem, cudaStream_t stream) {
    memset(cfg, 0, sizeof(*cfg));
    cfg->gridDim = grid; cfg->blockDim = block; cfg->dynamicSmemBytes = smem; cfg->stream = stream;
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = g_pdl ? 1 : 0;
    cfg->attrs = attr; cfg->numAttrs = 1;
}

// clusters / DSMEM
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r;
}
__device__ __forceinline__ uint32_t cluster_nctarank() {
    uint32_t r; asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r)); return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t dsmem_addr(const void* local_smem_ptr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_u32(local_smem_ptr)), "r"(rank));
    return r;
}
__device__ __forceinline__ void cluster_arrive_release() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_wait_acquire() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }
__device__ __forceinline__ void dsmem_st_f32x4(uint32_t addr, float a, float b, float c, float d) {
    asm volatile("st.shared::cluster.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
__device__ __forceinline__ float dsmem_ld_f32(uint32_t addr) {
    float v; asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(addr) : "memory"); return v;
}
__device__ __forceinline__ float4 dsmem_ld_f32x4(uint32_t addr) {
    float4 v;
    asm volatile("ld.shared::cluster.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr) : "memory");
    return v;
}

// tcgen05 / TMEM
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "elect.sync _|p, 0xffffffff;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(pred));
    return pred != 0;
}

// ---------------------------------------------------------------------------------------------
// optional in-stream event profiling of kernel categories (bench.py roofline section)
// ---------------------------------------------------------------------------------------------
enum { SAT_PROF_ATTN_FWD = 0, SAT_PROF_ATTN_BWD = 1, SAT_PROF_GEMM = 2, SAT_PROF_NCAT = 3 };
struct SatProfScope {
    int cat; cudaStream_t st; int slot;
    SatProfScope(int cat, cudaStream_t st);
    ~SatProfScope();
};
