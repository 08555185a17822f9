// Streaming (persistent, warp-specialised) kernels of the fused attention decode step -- the fast path for the
// shapes the configs use (D in {512, 2048}, H in {512, 768}).  sat_attn.cuh keeps the generic cluster kernels.
//
// WORK PARTITION ("stream-L").  The step's input is one long stream of slots: item 0's locations, item 1's ...
// (item = one image, or one image x one group of <= 5 beam rows; slot = LS consecutive locations = their Wa rows
// followed by their feature rows).  The stream is cut into G contiguous ranges, one per persistent CTA (G <= #SMs).
// Large batches are cut at item boundaries (nothing to combine); smaller ones are cut into equal slot counts so every
// SM moves the same number of bytes whatever the batch size -- an item that straddles a cut leaves one partial record
// per piece in a scratch buffer and a tiny second kernel (one block per cut item) combines them.  No atomics, no
// spinning, no co-residency requirement.
//
// INSIDE A CTA.  Warps 8-11 are PRODUCERS: one elected lane each keeps its share of a ring of shared-memory slots full
// with cp.async.bulk (TMA engine, mbarrier complete_tx), running ahead across item boundaries so HBM requests never
// pause (four issuing threads because one thread can only start a bulk copy every ~240 ns); producer 0 also fetches
// each item's small per-item vectors ("side block") so the consumers never wait on a global load.  Warps 0-7 are
// CONSUMERS (registers rebalanced with setmaxnreg): local slot q belongs to warp q % 8, which keeps a private
// online-softmax state (running max, running sum, un-normalised context accumulator of D/32 fp32 registers per lane
// and beam row) -- no block-wide synchronisation while data streams; per item the eight states are merged through
// shared memory with three named barriers.
// Each element of `a` and `Wa` crosses HBM->SM exactly once per step: L*(H+D)*sizeof(T) bytes per image.
#pragma once
#include "sat_attn.cuh"

extern int g_attn_generic;     // debug knob (sat_set_attn_generic): 1 = always use the generic cluster kernels
extern int g_attn_wa_hint_fwd, g_attn_wa_hint_bwd;   // L2 policy of the Wa rows, see Part
extern long long g_attn_pin_bytes;   // bytes of a / Wa per launch loaded with the L2 evict_last hint
extern int g_attn_pf_fwd, g_attn_pf_bwd;   // tuning knob (sat_set_attn_prefetch): L2 prefetch distance of the producers, in slots per CTA
extern int g_attn_part;        // debug knob (sat_set_attn_partition): 0 = auto, 1 = whole items, 2 = equal slot counts

namespace attn_s {
using attn::FwdParams;
using attn::BwdParams;

constexpr int CWARPS = 8;
constexpr int CTHREADS = CWARPS * 32;
constexpr int PWARPS = 4;                      // service warpgroup (setmaxnreg works per warpgroup): one producer lane per warp
constexpr int THREADS = CTHREADS + 32 * PWARPS;
constexpr int REGS_CONSUMER = 224, REGS_PRODUCER = 56;     // 256*224 + 128*56 = 384*168
constexpr int MAX_SLOTS = 32;
constexpr int REC_HDR = 16;                     // floats at the head of a partial record
constexpr size_t SMEM_MAX = 227 * 1024;

__device__ __forceinline__ void consumer_bar() { asm volatile("bar.sync 1, %0;" ::"n"(CTHREADS) : "memory"); }
template <int N> __device__ __forceinline__ void reg_inc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(N)); }
template <int N> __device__ __forceinline__ void reg_dec() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(N)); }
__device__ __forceinline__ void fence_acq_rel_gpu() { asm volatile("fence.acq_rel.gpu;" ::: "memory"); }

constexpr int pow2floor(int x) { int r = 1; while (2 * r <= x) r *= 2; return r; }
constexpr int clampi(int x, int lo, int hi) { return x < lo ? lo : (x > hi ? hi : x); }

template <typename T, int D, int H> struct Geo {
    static constexpr int VEC = Vec16<T>::N;
    static constexpr int DPL = D / VEC / 32;      // 16-byte chunks per lane per feature row
    static constexpr int HPL = H / VEC / 32;      // 16-byte chunks per lane per projection row
    static constexpr int ROWB = (H + D) * (int)sizeof(T);
    static constexpr int LS = clampi(pow2floor(12288 / ROWB), 1, 8);      // locations per slot
    static constexpr int SLOT = LS * ROWB;
    static_assert(D % (VEC * 32) == 0 && H % (VEC * 32) == 0, "D and H must be multiples of 32 16-byte chunks");
};

template <typename T> __device__ __forceinline__ uint4 lds16(const T* p) { return *reinterpret_cast<const uint4*>(p); }
__device__ __forceinline__ float4 ldcg4(const float* p) { return __ldcg(reinterpret_cast<const float4*>(p)); }

// contiguous cut of `tot` slots into G ranges whose boundaries are multiples of `align` slots
// (align = 1: equal slot counts, items may be cut;  align = slots per item: whole items only)
// the first `pin` slots of every item are loaded with the L2 evict_last hint (kept for the next step's kernels, which
// read the same a / Wa), the rest with evict_first -- the same fraction of every CTA's range whatever the partition
// wa_hint: L2 policy of the Wa rows (0 = like the rest of the slot, 1 = evict_last, 2 = evict_normal)
struct Part { long long tot; int align; int pin; int wa_hint; int pf; };     // pf: L2 prefetch distance in slots (0 = off)
__host__ __device__ __forceinline__ long long pstart(const Part& pt, int G, int i) { return (pt.tot / pt.align) * i / G * pt.align; }
__device__ __forceinline__ int cta_of(const Part& pt, int G, long long slot) {
    int i = (int)(slot * G / pt.tot);
    if (i > G - 1) i = G - 1;
    while (i + 1 < G && pstart(pt, G, i + 1) <= slot) ++i;
    while (i > 0 && pstart(pt, G, i) > slot) --i;
    return i;
}

struct Smem { size_t ring, red, es, wms, side, misc, bars, total; int nslots, lpad; };
constexpr int BAR_BYTES = (2 * MAX_SLOTS + 4) * 8;      // full[], empty[], side_full[2], side_empty[2]

inline Smem make_smem(int slot_bytes, size_t red_bytes, size_t es_bytes, size_t wms_bytes, size_t pcs_bytes, size_t side_bytes) {
    Smem s;
    const size_t misc = pcs_bytes;
    const size_t fixed = red_bytes + es_bytes + wms_bytes + side_bytes + misc + BAR_BYTES + 512;
    int ns = (int)((SMEM_MAX - fixed) / slot_bytes);
    if (ns > MAX_SLOTS) ns = MAX_SLOTS;
    s.nslots = ns;
    size_t o = 0;
    s.ring = o; o += (size_t)(ns > 0 ? ns : 0) * slot_bytes;
    o = (o + 127) & ~(size_t)127;
    s.red = o; o += red_bytes;
    s.es = o; o += es_bytes;
    s.wms = o; o += wms_bytes;
    o = (o + 127) & ~(size_t)127;
    s.side = o; o += side_bytes;
    o = (o + 15) & ~(size_t)15;
    s.misc = o; o += misc;
    o = (o + 15) & ~(size_t)15;
    s.bars = o; o += BAR_BYTES;
    s.total = o;
    s.lpad = 0;
    return s;
}

// producers: one elected lane in each of the PWARPS service warps.  A single thread can issue a bulk copy only every
// ~240 ns (measured: tools/tma_bw.cu -- 43 GB/s per SM with one issuing thread and 10 KB slots, 140-180 GB/s with 4-8), so
// the CTA's slot range [q0, q1) is dealt round-robin: producer p streams local slots n = p, p + PWARPS, ... through the ring.
// Producer 0 also fetches every segment's per-item vectors ("side block", double buffered) with
// side(img, grp, buffer, barrier) -- one thread, in segment order, so its parity waits can never run two phases ahead.
//
// Ring slot ownership: local slot n goes to consumer warp w = n % 8, which cycles through ITS ring slots
// {w, w+8, ...} < nslots, and is filled by producer w % PWARPS.  A ring slot (and its two mbarriers) therefore has exactly
// one producer and one consumer, neither of which can run a whole phase ahead of the barrier it waits on (the parity wait
// would otherwise pass on stale data).
template <typename T, int D, int H, typename SideFn>
__device__ __forceinline__ void produce(int pidx, bool static_inputs, const void* wa, const void* a, int L, int ngroups, long long q0, long long q1, int pin, int wa_hint, int pf, int spi,
                                        uint8_t* ring, uint64_t* full, uint64_t* empty, int nslots, uint64_t* sfull,
                                        uint64_t* sempty, SideFn side) {
    using G = Geo<T, D, H>;
    constexpr int LS = G::LS;
    static_assert(CWARPS == 2 * PWARPS, "each producer feeds two consumer warps");
    const long long item0 = q0 / spi;
    // position of this producer's next slot inside its item
    long long item = (q0 + pidx) / spi;
    int k = (int)(q0 + pidx - item * spi);
    long long img = item / ngroups;
    int grp = (int)(item - img * ngroups);
    // ring cursors of the two consumer warps this producer feeds: warp pidx (even visits) and warp pidx + PWARPS (odd visits)
    // (cursor "c" is the one used by the current slot, "n" by the next; swapped every slot)
    uint32_t rc = (uint32_t)pidx, rn = (uint32_t)(pidx + PWARPS), phc = 0, phn = 0, basec = rc, basen = rn;
    uint32_t endc = basec + CWARPS * (((uint32_t)nslots - basec + CWARPS - 1) >> 3);
    uint32_t endn = basen + CWARPS * (((uint32_t)nslots - basen + CWARPS - 1) >> 3);
    // side blocks (producer 0): segment `seg` starts at slot side_next
    uint32_t seg = 0;
    long long side_next = q0;
    auto side_upto = [&](long long q) {
        while (side_next <= q && side_next < q1) {
            const long long it = side_next / spi;
            const long long im = it / ngroups;
            const uint32_t sb = seg & 1;
            mbar_wait(&sempty[sb], ((seg >> 1) & 1) ^ 1);
            side(im, (int)(it - im * ngroups), (int)sb, &sfull[sb]);
            ++seg;
            side_next = (it + 1) * spi;
        }
    };
    (void)item0;
    // Programmatic dependent launch: this kernel may have started while the previous one is still running.  The side
    // blocks are the previous kernel's output and must wait for it; a / wa are older when the caller says so, and then
    // the first ring-full of slots is requested right away (producer 0: as many slots as it can issue without needing a
    // consumer, then the wait, then the side blocks the consumers are blocked on).
    bool waited = !static_inputs;
    if (!static_inputs) pdl_wait();
    const int prefill = (int)((endc - basec) >> 3) + (int)((endn - basen) >> 3);
    // L2 prefetch-ahead.  The ring (all the shared memory there is: ~16 slots = 160 KB per SM) bounds the bytes a CTA has in
    // flight, and with it the per-SM stream rate at HBM latency (ncu r2: both streaming kernels pull ~4 TB/s from DRAM while
    // their consumers wait for data 60 % of the time).  Requests that only have to reach L2 cost no shared memory: every
    // producer asks L2 for the slot it will load `pf` slots from now, so the later cp.async.bulk finds its lines in L2
    // (or already on their way) and the ring turns over at L2 latency instead of DRAM latency.
    const int pfd = (pf / PWARPS) * PWARPS;          // keep the prefetched slot in this producer's own sequence
    auto prefetch_slot = [&](long long q) {
        const long long it = q / spi;
        const int kk = (int)(q - it * spi);
        const long long im = it / ngroups;
        if (ngroups > 1 && it - im * ngroups != 0) return;          // beam groups re-read the image: prefetch it once
        const int l0 = kk * LS, nloc = min(LS, L - l0);
        l2_prefetch_bulk(reinterpret_cast<const T*>(wa) + ((size_t)im * L + l0) * H, (uint32_t)(nloc * H * sizeof(T)));
        l2_prefetch_bulk(reinterpret_cast<const T*>(a) + ((size_t)im * L + l0) * D, (uint32_t)(nloc * D * sizeof(T)));
    };
    if (pfd > 0 && static_inputs)
        for (long long q = q0 + pidx + prefill * PWARPS; q < q1 && q < q0 + pidx + pfd; q += PWARPS) prefetch_slot(q);
    int issued = 0;
    for (long long q = q0 + pidx; q < q1; q += PWARPS, ++issued) {
        if (pidx == 0 && (waited || issued >= prefill)) {
            if (!waited) { pdl_wait(); waited = true; }
            side_upto(q);
        }
        const uint32_t rr = rc;
        mbar_wait(&empty[rr], phc ^ 1);
        const int l0 = k * LS, nloc = min(LS, L - l0);
        uint8_t* dst = ring + (size_t)rr * G::SLOT;
        mbar_expect_tx(&full[rr], (uint32_t)(nloc * G::ROWB));
        const uint64_t pol = k < pin ? L2_EVICT_LAST : L2_EVICT_FIRST;
        bulk_g2s_hint(dst, reinterpret_cast<const T*>(wa) + ((size_t)img * L + l0) * H, (uint32_t)(nloc * H * sizeof(T)), &full[rr],
                      wa_hint == 1 ? L2_EVICT_LAST : (wa_hint == 2 ? L2_EVICT_NORMAL : pol));
        bulk_g2s_hint(dst + LS * H * sizeof(T), reinterpret_cast<const T*>(a) + ((size_t)img * L + l0) * D, (uint32_t)(nloc * D * sizeof(T)), &full[rr], pol);
        if (pfd > 0 && q + pfd < q1 && (waited || static_inputs)) prefetch_slot(q + pfd);
        rc += CWARPS;
        if (rc >= endc) { rc = basec; phc ^= 1; }
        { uint32_t t; t = rc; rc = rn; rn = t; t = phc; phc = phn; phn = t; t = basec; basec = basen; basen = t; t = endc; endc = endn; endn = t; }
        k += PWARPS;
        while (k >= spi) { k -= spi; if (++grp == ngroups) { grp = 0; ++img; } }
    }
    if (pidx == 0) {
        if (!waited) pdl_wait();
        side_upto(q1 - 1);
    }
}

// ---- items cut by the partition ----------------------------------------------------------------------------------
// A CTA range has at most two cut items: the one it starts inside (record 0) and the one it ends inside (record 1).
// The consumers merge their eight states and park the partial result in the CTA's record; the combine kernels below
// (one block per cut item, launched right after the streaming kernel) turn the pieces' records into the outputs.
struct Pieces { int ifirst, np; long long s0; };
__device__ __forceinline__ Pieces pieces_of(long long item, const Part& pt, int G, int spi) {
    Pieces pc;
    pc.s0 = item * spi;
    pc.ifirst = cta_of(pt, G, pc.s0);
    pc.np = cta_of(pt, G, pc.s0 + spi - 1) - pc.ifirst + 1;
    return pc;
}
// record slot of piece `i` of an item
__device__ __forceinline__ int piece_rec(const Pieces& pc, int i, const Part& pt, int G) {
    return (pc.ifirst + i) * 2 + ((i == 0 && pstart(pt, G, pc.ifirst) < pc.s0) ? 1 : 0);
}
// block `blockIdx.x` of a combine kernel looks at cut boundary blockIdx.x + 1; it owns the item containing that
// boundary iff the boundary is strictly inside the item and is the first such boundary of the item
__device__ __forceinline__ bool combine_owner(const Part& pt, int G, int spi, long long* item, Pieces* pc) {
    const int i = blockIdx.x + 1;
    const long long s = pstart(pt, G, i);
    const long long it = s / spi;
    if (s == it * spi || s >= pt.tot) return false;
    *pc = pieces_of(it, pt, G, spi);
    *item = it;
    return pc->ifirst + 1 == i;
}

// ------------------------------------------------------------------------------------------------------------
// forward
// ------------------------------------------------------------------------------------------------------------
template <typename T, int D, int H, int NBG>
inline Smem fwd_smem(int L) {
    const int lpad = (L + 3) & ~3;
    Smem s = make_smem(Geo<T, D, H>::SLOT, (size_t)4 * NBG * D * 4, (size_t)2 * NBG * lpad * 4, (size_t)2 * CWARPS * NBG * 2 * 4,
                       0, (size_t)2 * NBG * (H + D) * 4);
    s.lpad = lpad;
    return s;
}

// final write of 4 context columns of one row: fp32 copy for the backward pass, T copy, gated T copy (decoder.py:99-100)
template <typename T, bool PRECISE>
__device__ __forceinline__ void fwd_emit(const FwdParams& p, int64_t row, int d, float* c4, const float* gate_pre) {
    if (p.ctx32) *reinterpret_cast<float4*>(p.ctx32 + row * p.ld_c32 + d) = make_float4(c4[0], c4[1], c4[2], c4[3]);
    if (p.ctx) store4<T>(reinterpret_cast<T*>(p.ctx) + row * p.ld_ctx + d, c4);
    if (p.z) {
        if (p.gate) {
            const float4 gp = *reinterpret_cast<const float4*>(gate_pre);
            c4[0] *= sat_sigmoid<PRECISE>(gp.x); c4[1] *= sat_sigmoid<PRECISE>(gp.y);
            c4[2] *= sat_sigmoid<PRECISE>(gp.z); c4[3] *= sat_sigmoid<PRECISE>(gp.w);
        }
        store4<T>(reinterpret_cast<T*>(p.z) + row * p.ld_z + d, c4);
    }
}

template <typename T, int D, int H, int NBG, bool PRECISE>
__global__ void __launch_bounds__(THREADS, 1)
fwd_stream_kernel(FwdParams p, Smem sl, Part pt, int ngroups) {
    using G = Geo<T, D, H>;
    constexpr int VEC = G::VEC, DPL = G::DPL, HPL = G::HPL, LS = G::LS;
    constexpr int RECF = REC_HDR + NBG * D;
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t* ring = smem + sl.ring;
    float* red = reinterpret_cast<float*>(smem + sl.red);          // [4][NBG][D]
    float* e_s = reinterpret_cast<float*>(smem + sl.es);           // [2][NBG][lpad]
    float* wms = reinterpret_cast<float*>(smem + sl.wms);          // [2][CWARPS][NBG][2]
    float* side = reinterpret_cast<float*>(smem + sl.side);        // [2][NBG][H + D]: the item's [U h | f_beta h] rows
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + sl.bars);
    uint64_t* empty = full + MAX_SLOTS;
    uint64_t* sfull = empty + MAX_SLOTS;
    uint64_t* sempty = sfull + 2;
    constexpr int SIDE_F = NBG * (H + D);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int L = p.L, nslots = sl.nslots, lpad = sl.lpad, Gn = gridDim.x;
    const int spi = (L + LS - 1) / LS;                              // slots per item
    const long long q0 = pstart(pt, Gn, blockIdx.x), q1 = pstart(pt, Gn, blockIdx.x + 1);

    if (tid == 0) {
        for (int i = 0; i < nslots; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(&sfull[i], 1); mbar_init(&sempty[i], CWARPS); }
        mbar_fence_init();
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();

    if (warp < CWARPS) {
        reg_inc<REGS_CONSUMER>();
        pdl_wait();                                   // everything this kernel reads besides a / wa, and all it writes
        if (tid == 0) pdl_launch_dependents();

        float vreg[HPL * VEC];
#pragma unroll
        for (int j = 0; j < HPL; ++j)
#pragma unroll
            for (int q = 0; q < VEC / 4; ++q) {
                const float4 t = *reinterpret_cast<const float4*>(p.v + (lane + 32 * j) * VEC + 4 * q);
                vreg[j * VEC + 4 * q] = t.x; vreg[j * VEC + 4 * q + 1] = t.y; vreg[j * VEC + 4 * q + 2] = t.z; vreg[j * VEC + 4 * q + 3] = t.w;
            }
        const float bv = p.bv[0];
        long long ql = warp;                 // this warp's next local slot
        const uint32_t nw = ((uint32_t)nslots - warp + CWARPS - 1) >> 3;   // ring slots owned by this warp: warp, warp+8, ...
        uint32_t r = warp, ph = 0;           // its ring position
        int par = 0;
        uint32_t seg = 0;
        for (long long gs = q0; gs < q1; par ^= 1, ++seg) {
            const long long item = gs / spi;
            const long long ge = min(q1, (item + 1) * spi);
            const int k0 = (int)(gs - item * spi), k1 = (int)(ge - item * spi);
            const bool whole = (k0 == 0 && k1 == spi);
            const int img = (int)(item / ngroups), grp = (int)(item - (long long)img * ngroups);
            const int row_base = img * p.NB + grp * NBG;
            const int nb_valid = min(NBG, p.NB - grp * NBG);
            const float* uf_s = side + (size_t)(seg & 1) * SIDE_F;     // this item's uf rows, fetched by the producer
            mbar_wait(&sfull[seg & 1], (seg >> 1) & 1);
            // fp32 mode with beam rows: tanh(Wa + Uh) from separate exponentials (see tanh_from_exp2x).  The consumers turn the
            // item's Uh rows into e^{2 Uh} in place, once, before anyone reads them.
            constexpr bool EXPU = PRECISE && NBG > 1;
            if constexpr (EXPU) {
                float* us = side + (size_t)(seg & 1) * SIDE_F;
                for (int i = tid; i < NBG * H; i += CTHREADS) {
                    const int nb = i / H, hh = i - nb * H;
                    us[nb * (H + D) + hh] = exp2x_clamped(us[nb * (H + D) + hh]);
                }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");      // these generic writes precede the producer's next bulk copy
                consumer_bar();
            }
            float uh[HPL * VEC];
            if (NBG == 1) {
#pragma unroll
                for (int j = 0; j < HPL; ++j)
#pragma unroll
                    for (int q = 0; q < VEC / 4; ++q) {
                        const float4 t = *reinterpret_cast<const float4*>(uf_s + (lane + 32 * j) * VEC + 4 * q);
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
            const long long seg_end_local = ge - q0;

            for (; ql < seg_end_local; ql += CWARPS) {
                const int k = (int)(q0 + ql - item * spi);
                mbar_wait(&full[r], ph);
                const T* wa_s = reinterpret_cast<const T*>(ring + (size_t)r * G::SLOT);
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
                        if constexpr (EXPU) {
#pragma unroll
                            for (int i = 0; i < VEC; ++i) { f0[i] = exp2x_clamped(f0[i]); f1[i] = exp2x_clamped(f1[i]); }
                        }
#pragma unroll
                        for (int nb = 0; nb < NBG; ++nb) {
                            float u[VEC];
                            if (NBG == 1) {
#pragma unroll
                                for (int i = 0; i < VEC; ++i) u[i] = uh[j * VEC + i];
                            } else {
#pragma unroll
                                for (int q = 0; q < VEC / 4; ++q) {
                                    const float4 t = *reinterpret_cast<const float4*>(uf_s + nb * (H + D) + c * VEC + 4 * q);
                                    u[4 * q] = t.x; u[4 * q + 1] = t.y; u[4 * q + 2] = t.z; u[4 * q + 3] = t.w;
                                }
                            }
#pragma unroll
                            for (int i = 0; i < VEC; ++i) {
                                if constexpr (EXPU) {                       // f0 / f1 hold e^{2 Wa}, u holds e^{2 Uh}
                                    sc0[nb] = fmaf(vreg[j * VEC + i], tanh_from_exp2x(f0[i] * u[i]), sc0[nb]);
                                    sc1[nb] = fmaf(vreg[j * VEC + i], tanh_from_exp2x(f1[i] * u[i]), sc1[nb]);
                                } else {
                                    sc0[nb] = fmaf(vreg[j * VEC + i], attn_tanh<PRECISE>(f0[i] + u[i]), sc0[nb]);
                                    sc1[nb] = fmaf(vreg[j * VEC + i], attn_tanh<PRECISE>(f1[i] + u[i]), sc1[nb]);
                                }
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
                            const float rs = sat_exp<PRECISE>(m[nb] - mn);
                            s[nb] *= rs;
#pragma unroll
                            for (int i = 0; i < DPL * VEC; ++i) acc[nb][i] *= rs;
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
                if (lane == 0) mbar_arrive(&empty[r]);
                r += CWARPS;
                if (r >= (uint32_t)warp + CWARPS * nw) { r = warp; ph ^= 1; }
            }

            // ---- merge the eight per-warp online-softmax states of this segment ----
            float* wms_item = wms + (size_t)par * CWARPS * NBG * 2;
            if (lane == 0) {
#pragma unroll
                for (int nb = 0; nb < NBG; ++nb) { wms_item[(warp * NBG + nb) * 2] = m[nb]; wms_item[(warp * NBG + nb) * 2 + 1] = s[nb]; }
            }
            consumer_bar();                                                   // B1: all (m, s) and all scores are published
            float Mx[NBG], Sx[NBG];
#pragma unroll
            for (int nb = 0; nb < NBG; ++nb) {
                float M = -INFINITY;
#pragma unroll
                for (int w = 0; w < CWARPS; ++w) M = fmaxf(M, wms_item[(w * NBG + nb) * 2]);
                float S = 0.0f;
#pragma unroll
                for (int w = 0; w < CWARPS; ++w) S = fmaf(wms_item[(w * NBG + nb) * 2 + 1], sat_exp<PRECISE>(wms_item[(w * NBG + nb) * 2] - M), S);
                Mx[nb] = M; Sx[nb] = S;
                const float f = sat_exp<PRECISE>(m[nb] - M);
#pragma unroll
                for (int i = 0; i < DPL * VEC; ++i) acc[nb][i] *= f;
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
            consumer_bar();                                                   // B3: red[0..3] hold the CTA's partial context

            // selects of the small per-row register arrays by a runtime row index
            auto pick = [&](const float* arr, int nb) { float x = arr[0];
#pragma unroll
                for (int k = 1; k < NBG; ++k) if (nb == k) x = arr[k];
                return x; };
            if (whole) {
                float inv[NBG];
#pragma unroll
                for (int nb = 0; nb < NBG; ++nb) inv[nb] = 1.0f / Sx[nb];
                for (int i = tid; i < nb_valid * L; i += CTHREADS) {          // alphas (fp32, decoder.py:128)
                    const int nb = i / L, l = i - nb * L;
                    p.alpha[(int64_t)(row_base + nb) * p.ld_alpha + l] = sat_exp<PRECISE>(e_item[nb * lpad + l] - pick(Mx, nb)) * pick(inv, nb);
                }
                for (int i = tid; i < nb_valid * (D / 4); i += CTHREADS) {
                    const int nb = i / (D / 4), d = 4 * (i - nb * (D / 4));
                    const float iv = pick(inv, nb);
                    float4 t = *reinterpret_cast<const float4*>(red + (size_t)nb * D + d);
#pragma unroll
                    for (int w = 1; w < 4; ++w) {
                        const float4 x = *reinterpret_cast<const float4*>(red + (size_t)(w * NBG + nb) * D + d);
                        t.x += x.x; t.y += x.y; t.z += x.z; t.w += x.w;
                    }
                    float c4[4] = {t.x * iv, t.y * iv, t.z * iv, t.w * iv};
                    fwd_emit<T, PRECISE>(p, row_base + nb, d, c4, uf_s + nb * (H + D) + H + d);
                }
            } else {
                // ---- this CTA holds only the locations [lb, le) of the item: publish a partial record ----
                const int lb = k0 * LS, le = min(L, k1 * LS);
                float* rec = p.s_part + (size_t)(blockIdx.x * 2 + (gs == q0 ? 0 : 1)) * RECF;
                for (int i = tid; i < nb_valid * (le - lb); i += CTHREADS) {  // raw scores parked in the alpha rows
                    const int nb = i / (le - lb), l = lb + (i - nb * (le - lb));
                    p.alpha[(int64_t)(row_base + nb) * p.ld_alpha + l] = e_item[nb * lpad + l];
                }
                if (tid < NBG) { rec[tid * 2] = pick(Mx, tid); rec[tid * 2 + 1] = pick(Sx, tid); }
                for (int i = tid; i < NBG * (D / 4); i += CTHREADS) {
                    const int nb = i / (D / 4), d = 4 * (i - nb * (D / 4));
                    float4 t = *reinterpret_cast<const float4*>(red + (size_t)nb * D + d);
#pragma unroll
                    for (int w = 1; w < 4; ++w) {
                        const float4 x = *reinterpret_cast<const float4*>(red + (size_t)(w * NBG + nb) * D + d);
                        t.x += x.x; t.y += x.y; t.z += x.z; t.w += x.w;
                    }
                    *reinterpret_cast<float4*>(rec + REC_HDR + (size_t)nb * D + d) = t;
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&sempty[seg & 1]);          // this item's side block may be overwritten
            gs = ge;
        }
    } else {
        reg_dec<REGS_PRODUCER>();
        if (lane == 0) {
            auto side_fn = [&](long long img, int grp, int sb, uint64_t* bar) {
                const int nbv = min(NBG, p.NB - grp * NBG);
                mbar_expect_tx(bar, (uint32_t)(NBG * (H + D) * 4));
                for (int nb = 0; nb < NBG; ++nb) {          // rows beyond the live beam repeat the last valid row
                    const int64_t row = img * p.NB + grp * NBG + min(nb, nbv - 1);
                    bulk_g2s(side + (size_t)sb * SIDE_F + (size_t)nb * (H + D), p.uf + row * p.ld_uf, (uint32_t)((H + D) * 4), bar);
                }
            };
            produce<T, D, H>(warp - CWARPS, p.static_inputs != 0, p.wa, p.a, L, ngroups, q0, q1, pt.pin, pt.wa_hint, pt.pf, spi, ring, full, empty, nslots, sfull, sempty, side_fn);
        }
    }
}

// combine kernel of the forward step: one block per item the partition cut (see combine_owner)
template <typename T, int D, int H, int NBG, bool PRECISE>
__global__ void __launch_bounds__(256)
fwd_combine_kernel(FwdParams p, Part pt, int Gn, int ngroups) {
    using G = Geo<T, D, H>;
    constexpr int RECF = REC_HDR + NBG * D;
    const int spi = (p.L + G::LS - 1) / G::LS;
    long long item; Pieces pc;
    if (!combine_owner(pt, Gn, spi, &item, &pc)) return;
    const int tid = threadIdx.x, L = p.L;
    const int img = (int)(item / ngroups), grp = (int)(item - (long long)img * ngroups);
    const int row_base = img * p.NB + grp * NBG;
    const int nb_valid = min(NBG, p.NB - grp * NBG);
    __shared__ float Ms[NBG], Is[NBG];
    __shared__ int recs[64];
    for (int q = tid; q < pc.np && q < 64; q += blockDim.x) recs[q] = piece_rec(pc, q, pt, Gn);
    __syncthreads();
    if (tid < nb_valid) {
        float M = -INFINITY, S = 0.0f;
        for (int q = 0; q < pc.np; ++q) M = fmaxf(M, p.s_part[(size_t)recs[q] * RECF + tid * 2]);
        for (int q = 0; q < pc.np; ++q) {
            const float* r = p.s_part + (size_t)recs[q] * RECF + tid * 2;
            S = fmaf(r[1], sat_exp<PRECISE>(r[0] - M), S);
        }
        Ms[tid] = M; Is[tid] = 1.0f / S;
    }
    __syncthreads();
    for (int i = tid; i < nb_valid * L; i += blockDim.x) {                 // raw score -> alpha
        const int nb = i / L, l = i - nb * L;
        float* ap = p.alpha + (int64_t)(row_base + nb) * p.ld_alpha + l;
        *ap = sat_exp<PRECISE>(*ap - Ms[nb]) * Is[nb];
    }
    for (int i = tid; i < nb_valid * (D / 4); i += blockDim.x) {
        const int nb = i / (D / 4), d = 4 * (i - nb * (D / 4));
        float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int q = 0; q < pc.np; ++q) {
            const float* r = p.s_part + (size_t)recs[q] * RECF;
            const float f = sat_exp<PRECISE>(r[nb * 2] - Ms[nb]);
            const float4 x = *reinterpret_cast<const float4*>(r + REC_HDR + (size_t)nb * D + d);
            t.x = fmaf(f, x.x, t.x); t.y = fmaf(f, x.y, t.y); t.z = fmaf(f, x.z, t.z); t.w = fmaf(f, x.w, t.w);
        }
        const float iv = Is[nb];
        float c4[4] = {t.x * iv, t.y * iv, t.z * iv, t.w * iv};
        const int64_t row = row_base + nb;
        fwd_emit<T, PRECISE>(p, row, d, c4, p.uf + row * p.ld_uf + H + d);
    }
}

// ------------------------------------------------------------------------------------------------------------
// backward of one attention step (NB == 1): the same stream, one pass.
// The softmax backward needs S = sum_l alpha_l dalpha_l before any de_l; with dalpha_l = dctx . a_l + galpha_l this
// is S = dctx . ctx + sum_l alpha_l galpha_l, and ctx is the (fp32) context the forward kernel saved -- so every
// location is finished the moment its two rows arrive:
//   dalpha_l = dctx . a_l (+ galpha_l);  de_l = alpha_l (dalpha_l - S);  th = tanh(Wa_l + Uh);
//   dpre_l = de_l v (1 - th^2)  -> written ONCE, in T, to this step's slice of the (T,B,L,H) dpre buffer (the sum
//   over time is one pass at the end of BPTT instead of a fp32 read-modify-write per step);
//   per warp: dUh += dpre_l, dv += de_l th, sum de_l; merged per item through shared memory / partial records.
// HBM traffic per image per step: L (H+D) sizeof(T) read + L H sizeof(T) written.
// ------------------------------------------------------------------------------------------------------------
template <typename T, int D, int H>
inline Smem bwd_smem(int L) {
    const int lpad = (L + 3) & ~3;
    // side block per buffer: [dctx D | ctx32 D | Uh H] fp32 (producer, TMA) + [alpha lpad | galpha lpad] (consumers)
    Smem s = make_smem(Geo<T, D, H>::SLOT, (size_t)CWARPS * 2 * H * 4, (size_t)2 * 2 * lpad * 4, (size_t)2 * CWARPS * 4, 0,
                       (size_t)2 * (2 * D + H) * 4);
    s.lpad = lpad;
    return s;
}
constexpr int BWD_MAXL = 1024;       // alpha / galpha rows are staged through 4 registers per consumer thread

// dctx = dz * sigmoid(f_beta pre) (+ extra);  d f_beta pre = dz * ctx * g (1 - g)   (backward of decoder.py:99-100).
// Used when the caller hands the un-gated upstream gradient to sat_attn_step_bwd; the decoder driver fuses this into the
// epilogue of the GEMM that produces dz instead (EpiDzGate).
template <typename T, bool PRECISE>
__global__ void gate_bwd_kernel(const float* __restrict__ dz, int64_t ld_dz, const float* __restrict__ uf, int64_t ld_uf,
                                const float* __restrict__ extra, int64_t ld_ex, const float* __restrict__ ctx32, int64_t ld_c32,
                                float* __restrict__ dctx, T* __restrict__ duf, int64_t ld_duf, int B, int D, int H, int gate) {
    const int b = blockIdx.y;
    for (int d = blockIdx.x * blockDim.x + threadIdx.x; d < D; d += gridDim.x * blockDim.x) {
        const float z = dz[(int64_t)b * ld_dz + d];
        float g = 1.0f, dfb = 0.0f;
        if (gate) {
            g = sat_sigmoid<PRECISE>(uf[(int64_t)b * ld_uf + H + d]);
            dfb = z * ctx32[(int64_t)b * ld_c32 + d] * g * (1.0f - g);
        }
        dctx[(size_t)b * D + d] = z * g + (extra ? extra[(int64_t)b * ld_ex + d] : 0.0f);
        duf[(int64_t)b * ld_duf + H + d] = from_f32<T>(dfb);
    }
}

// p.dz holds dctx (fp32, row pitch ld_dz) -- see gate_bwd_kernel / EpiDzGate; p.gate / p.dctx_extra are not used here.
template <typename T, int D, int H, bool PRECISE>
__global__ void __launch_bounds__(THREADS, 1)
bwd_stream_kernel(BwdParams p, Smem sl, Part pt) {
    using G = Geo<T, D, H>;
    constexpr int VEC = G::VEC, DPL = G::DPL, HPL = G::HPL, LS = G::LS;
    constexpr int RECF = REC_HDR + 2 * H;
    constexpr int SIDE_F = 2 * D + H;
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t* ring = smem + sl.ring;
    float* red = reinterpret_cast<float*>(smem + sl.red);          // [CWARPS][2][H]
    float* al_s = reinterpret_cast<float*>(smem + sl.es);          // [2][2][lpad]: alpha row, galpha row
    float* wsum = reinterpret_cast<float*>(smem + sl.wms);         // [2][CWARPS]
    float* side = reinterpret_cast<float*>(smem + sl.side);        // [2][dctx D | ctx32 D | Uh H]
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + sl.bars);
    uint64_t* empty = full + MAX_SLOTS;
    uint64_t* sfull = empty + MAX_SLOTS;
    uint64_t* sempty = sfull + 2;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int L = p.L, nslots = sl.nslots, lpad = sl.lpad, Gn = gridDim.x;
    const int spi = (L + LS - 1) / LS;
    const long long q0 = pstart(pt, Gn, blockIdx.x), q1 = pstart(pt, Gn, blockIdx.x + 1);

    if (tid == 0) {
        for (int i = 0; i < nslots; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(&sfull[i], 1); mbar_init(&sempty[i], CWARPS); }
        mbar_fence_init();
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();

    if (warp < CWARPS) {
        reg_inc<REGS_CONSUMER>();
        pdl_wait();                                   // everything this kernel reads besides a / wa, and all it writes
        if (tid == 0) pdl_launch_dependents();

        float vreg[HPL * VEC];
#pragma unroll
        for (int j = 0; j < HPL; ++j)
#pragma unroll
            for (int q = 0; q < VEC / 4; ++q) {
                const float4 t = *reinterpret_cast<const float4*>(p.v + (lane + 32 * j) * VEC + 4 * q);
                vreg[j * VEC + 4 * q] = t.x; vreg[j * VEC + 4 * q + 1] = t.y; vreg[j * VEC + 4 * q + 2] = t.z; vreg[j * VEC + 4 * q + 3] = t.w;
            }
        // alpha / galpha rows of an item: staged global -> registers -> shared one segment ahead, so that no consumer ever
        // waits on a global load while the stream runs
        float pa[BWD_MAXL / CTHREADS], pg[BWD_MAXL / CTHREADS];
        auto al_fetch = [&](long long b) {
#pragma unroll
            for (int i = 0; i < BWD_MAXL / CTHREADS; ++i) {
                const int l = tid + i * CTHREADS;
                pa[i] = (l < L) ? p.alpha[b * p.ld_alpha + l] : 0.0f;
                pg[i] = (l < L && p.galpha) ? p.galpha[b * p.ld_ga + l] : 0.0f;
            }
        };
        auto al_store = [&](int buf) {
#pragma unroll
            for (int i = 0; i < BWD_MAXL / CTHREADS; ++i) {
                const int l = tid + i * CTHREADS;
                if (l < L) { al_s[(size_t)(buf * 2) * lpad + l] = pa[i]; al_s[(size_t)(buf * 2 + 1) * lpad + l] = pg[i]; }
            }
        };
        if (q0 < q1) { al_fetch(q0 / spi); al_store(0); }
        consumer_bar();

        long long ql = warp;
        const uint32_t nw = ((uint32_t)nslots - warp + CWARPS - 1) >> 3;
        uint32_t r = warp, ph = 0;
        int par = 0;
        uint32_t seg = 0;
        for (long long gs = q0; gs < q1; par ^= 1, ++seg) {
            const long long b = gs / spi;
            const long long ge = min(q1, (b + 1) * spi);
            const int k0 = (int)(gs - b * spi), k1 = (int)(ge - b * spi);
            const bool whole = (k0 == 0 && k1 == spi);
            const int sb = seg & 1;
            const float* alpha_s = al_s + (size_t)(sb * 2) * lpad;
            const float* galpha_s = alpha_s + lpad;
            if (ge < q1) al_fetch(b + 1);                                  // next segment's rows: in flight during this one

            // ---- per-image vectors from the side block: dctx (registers, this lane's columns), S, Uh ----
            const float* sd = side + (size_t)sb * SIDE_F;
            mbar_wait(&sfull[sb], (seg >> 1) & 1);
            float dctx[DPL * VEC];
            float sdot = 0.0f;
#pragma unroll
            for (int j = 0; j < DPL; ++j)
#pragma unroll
                for (int q = 0; q < VEC / 4; ++q) {
                    const int d = (lane + 32 * j) * VEC + 4 * q;
                    const float4 dc = *reinterpret_cast<const float4*>(sd + d);
                    const float4 cx = *reinterpret_cast<const float4*>(sd + D + d);
                    dctx[j * VEC + 4 * q] = dc.x; dctx[j * VEC + 4 * q + 1] = dc.y; dctx[j * VEC + 4 * q + 2] = dc.z; dctx[j * VEC + 4 * q + 3] = dc.w;
                    sdot = fmaf(dc.x, cx.x, fmaf(dc.y, cx.y, fmaf(dc.z, cx.z, fmaf(dc.w, cx.w, sdot))));
                }
            for (int l = lane; l < L; l += 32) sdot = fmaf(alpha_s[l], galpha_s[l], sdot);
            const float S = warp_sum(sdot);
            float uh[HPL * VEC], dU[HPL * VEC], dV[HPL * VEC];
#pragma unroll
            for (int j = 0; j < HPL; ++j)
#pragma unroll
                for (int q = 0; q < VEC / 4; ++q) {
                    const float4 t = *reinterpret_cast<const float4*>(sd + 2 * D + (lane + 32 * j) * VEC + 4 * q);
                    uh[j * VEC + 4 * q] = t.x; uh[j * VEC + 4 * q + 1] = t.y; uh[j * VEC + 4 * q + 2] = t.z; uh[j * VEC + 4 * q + 3] = t.w;
                }
            __syncwarp();
            if (lane == 0) mbar_arrive(&sempty[sb]);                       // the side block now lives in registers
#pragma unroll
            for (int i = 0; i < HPL * VEC; ++i) { dU[i] = 0.0f; dV[i] = 0.0f; }
            float desum = 0.0f;
            T* dpre_g = reinterpret_cast<T*>(p.dpre_out) + (size_t)b * L * H;
            const bool keep_dpre = p.dpre_out != nullptr;
            float* de_g = p.de_out ? p.de_out + (size_t)b * L : nullptr;
            const long long seg_end_local = ge - q0;

            for (; ql < seg_end_local; ql += CWARPS) {
                const int k = (int)(q0 + ql - b * spi);
                mbar_wait(&full[r], ph);
                const T* wa_s = reinterpret_cast<const T*>(ring + (size_t)r * G::SLOT);
                const T* a_s = wa_s + LS * H;
                const int l0 = k * LS, nloc = min(LS, L - l0);
#pragma unroll 1
                for (int lp = 0; lp < nloc; lp += 2) {
                    const bool two = lp + 1 < nloc;
                    const int lq = two ? lp + 1 : lp;
                    const int l = l0 + lp;
                    float ps0[4] = {0.0f, 0.0f, 0.0f, 0.0f}, ps1[4] = {0.0f, 0.0f, 0.0f, 0.0f};   // 8 independent FMA chains
#pragma unroll
                    for (int j = 0; j < DPL; ++j) {
                        const int c = lane + 32 * j;
                        float g0[VEC], g1[VEC];
                        Vec16<T>::unpack(lds16(a_s + (size_t)lp * D + c * VEC), g0);
                        Vec16<T>::unpack(lds16(a_s + (size_t)lq * D + c * VEC), g1);
#pragma unroll
                        for (int i = 0; i < VEC; ++i) {
                            ps0[i & 3] = fmaf(g0[i], dctx[j * VEC + i], ps0[i & 3]);
                            ps1[i & 3] = fmaf(g1[i], dctx[j * VEC + i], ps1[i & 3]);
                        }
                    }
                    float s0 = (ps0[0] + ps0[1]) + (ps0[2] + ps0[3]), s1 = (ps1[0] + ps1[1]) + (ps1[2] + ps1[3]);
                    s0 = warp_sum(s0) + galpha_s[l]; s1 = warp_sum(s1) + galpha_s[two ? l + 1 : l];
                    const float de0 = alpha_s[l] * (s0 - S);                       // softmax backward
                    const float de1 = two ? alpha_s[l + 1] * (s1 - S) : 0.0f;
                    desum += de0 + de1;
                    if (de_g && lane == 0) { de_g[l] = de0; if (two) de_g[l + 1] = de1; }
#pragma unroll
                    for (int j = 0; j < HPL; ++j) {
                        const int c = lane + 32 * j;
                        float w0[VEC], w1[VEC], dp0[VEC], dp1[VEC];
                        Vec16<T>::unpack(lds16(wa_s + (size_t)lp * H + c * VEC), w0);
                        Vec16<T>::unpack(lds16(wa_s + (size_t)lq * H + c * VEC), w1);
#pragma unroll
                        for (int i = 0; i < VEC; ++i) {
                            const float t0 = attn_tanh<PRECISE>(w0[i] + uh[j * VEC + i]);
                            const float t1 = attn_tanh<PRECISE>(w1[i] + uh[j * VEC + i]);
                            dp0[i] = de0 * vreg[j * VEC + i] * (1.0f - t0 * t0);
                            dp1[i] = de1 * vreg[j * VEC + i] * (1.0f - t1 * t1);
                            dU[j * VEC + i] += dp0[i] + dp1[i];
                            dV[j * VEC + i] = fmaf(de0, t0, fmaf(de1, t1, dV[j * VEC + i]));
                        }
                        if (keep_dpre) {
                            __stcs(reinterpret_cast<uint4*>(dpre_g + (size_t)l * H + c * VEC), Vec16<T>::pack(dp0));     // written once, read at
                            if (two) __stcs(reinterpret_cast<uint4*>(dpre_g + (size_t)(l + 1) * H + c * VEC), Vec16<T>::pack(dp1));   // the end of BPTT
                        }
                    }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&empty[r]);
                r += CWARPS;
                if (r >= (uint32_t)warp + CWARPS * nw) { r = warp; ph ^= 1; }
            }

            // ---- cross-warp reduction of dUh / dv / sum de ----
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
            float* ws_item = wsum + par * CWARPS;
            if (lane == 0) ws_item[warp] = desum;
            consumer_bar();
            auto emit = [&](int h, float du, float dv) {
                reinterpret_cast<T*>(p.duf)[b * p.ld_duf + h] = from_f32<T>(du);
                float* dst = p.dv_acc + (size_t)b * H + h;
                *dst = p.first ? dv : *dst + dv;
            };
            if (whole) {
                for (int h = tid; h < H; h += CTHREADS) {
                    float du = 0.0f, dv = 0.0f;
#pragma unroll
                    for (int w = 0; w < CWARPS; ++w) { du += red[(size_t)(w * 2) * H + h]; dv += red[(size_t)(w * 2 + 1) * H + h]; }
                    emit(h, du, dv);
                }
                if (tid == 0) {
                    float t = 0.0f;
#pragma unroll
                    for (int w = 0; w < CWARPS; ++w) t += ws_item[w];
                    p.dbv_acc[b] = p.first ? t : p.dbv_acc[b] + t;
                }
            } else {
                float* rec = p.s_part + (size_t)(blockIdx.x * 2 + (gs == q0 ? 0 : 1)) * RECF;
                for (int h = tid; h < H; h += CTHREADS) {
                    float du = 0.0f, dv = 0.0f;
#pragma unroll
                    for (int w = 0; w < CWARPS; ++w) { du += red[(size_t)(w * 2) * H + h]; dv += red[(size_t)(w * 2 + 1) * H + h]; }
                    rec[REC_HDR + h] = du; rec[REC_HDR + H + h] = dv;
                }
                if (tid == 0) {
                    float t = 0.0f;
#pragma unroll
                    for (int w = 0; w < CWARPS; ++w) t += ws_item[w];
                    rec[0] = t;
                }
            }
            if (ge < q1) al_store(sb ^ 1);
            consumer_bar();          // publishes the next segment's alpha rows; also `red` is single-buffered
            gs = ge;
        }
    } else {
        reg_dec<REGS_PRODUCER>();
        if (lane == 0) {
            auto side_fn = [&](long long b, int, int sb, uint64_t* bar) {
                float* dst = side + (size_t)sb * SIDE_F;
                mbar_expect_tx(bar, (uint32_t)(SIDE_F * 4));
                bulk_g2s(dst, p.dz + b * p.ld_dz, (uint32_t)(D * 4), bar);
                bulk_g2s(dst + D, p.ctx32 + b * p.ld_c32, (uint32_t)(D * 4), bar);
                bulk_g2s(dst + 2 * D, p.uf + b * p.ld_uf, (uint32_t)(H * 4), bar);
            };
            produce<T, D, H>(warp - CWARPS, p.static_inputs != 0, p.wa, p.a, L, 1, q0, q1, pt.pin, pt.wa_hint, pt.pf, spi, ring, full, empty, nslots, sfull, sempty, side_fn);
        }
    }
}

// combine kernel of the backward step: one block per image the partition cut
template <typename T, int D, int H>
__global__ void __launch_bounds__(256)
bwd_combine_kernel(BwdParams p, Part pt, int Gn) {
    using G = Geo<T, D, H>;
    constexpr int RECF = REC_HDR + 2 * H;
    const int spi = (p.L + G::LS - 1) / G::LS;
    long long b; Pieces pc;
    if (!combine_owner(pt, Gn, spi, &b, &pc)) return;
    const int tid = threadIdx.x;
    __shared__ int recs[64];
    for (int q = tid; q < pc.np && q < 64; q += blockDim.x) recs[q] = piece_rec(pc, q, pt, Gn);
    __syncthreads();
    for (int h = tid; h < H; h += blockDim.x) {
        float du = 0.0f, dv = 0.0f;
        for (int q = 0; q < pc.np; ++q) {
            const float* rp = p.s_part + (size_t)recs[q] * RECF + REC_HDR;
            du += rp[h]; dv += rp[H + h];
        }
        reinterpret_cast<T*>(p.duf)[b * p.ld_duf + h] = from_f32<T>(du);
        float* dst = p.dv_acc + (size_t)b * H + h;
        *dst = p.first ? dv : *dst + dv;
    }
    if (tid == 0) {
        float t = 0.0f;
        for (int q = 0; q < pc.np; ++q) t += p.s_part[(size_t)recs[q] * RECF];
        p.dbv_acc[b] = p.first ? t : p.dbv_acc[b] + t;
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

inline bool shape_ok(int L, int D, int H) { return !g_attn_generic && (D == 512 || D == 2048) && (H == 512 || H == 768) && L >= 1 && L <= BWD_MAXL; }
inline bool fwd_shape_ok(int NB, int L, int D, int H) { return shape_ok(L, D, H) && NB >= 1 && (NB == 1 || D == 512); }

// scratch = [partial records: 2 per CTA][NB == 1: (NI, D) fp32 dctx of the backward]; no initialisation needed
inline int64_t scratch_rec_bytes(int NB, int D, int H) {
    const int nbg = NB == 1 ? 1 : 5;
    int64_t rec = REC_HDR + (int64_t)nbg * D;
    if (rec < REC_HDR + 2 * H) rec = REC_HDR + 2 * H;
    return ((int64_t)num_sms() * 2 * rec * 4 + 255) & ~(int64_t)255;
}
inline int64_t scratch_bytes(int NI, int NB, int D, int H) {
    return scratch_rec_bytes(NB, D, H) + (NB == 1 ? (int64_t)NI * D * 4 : 0);
}

// partition + grid.  Whole items per CTA unless the batch is tiny: measured on B200 (tools/attn_bench.cu) the uneven item
// counts cost less than the combine launch from ~1/4 of the SMs upwards, and from two items per SM the step is HBM-bound
// anyway.  Tiny batches (single-image beam search) get equal slot counts, never a range shorter than 1/8 of an item
// (bounds the pieces per item).
inline Part pick_part(long long nitems, int spi, int slot_bytes, int ngroups, bool bwd, int* grid, int max_ctas = 0) {
    int sms = num_sms();
    const bool capped = max_ctas > 0 && max_ctas < sms;
    if (capped) sms = max_ctas;
    Part pt;
    pt.tot = nitems * spi;
    // beam groups of one image re-read the image inside the launch: count each image's bytes once
    const double total_bytes = (double)pt.tot * slot_bytes / ngroups;
    pt.wa_hint = bwd ? g_attn_wa_hint_bwd : g_attn_wa_hint_fwd;
    pt.pf = bwd ? g_attn_pf_bwd : g_attn_pf_fwd;
    pt.pin = (double)g_attn_pin_bytes >= total_bytes ? spi : (int)(spi * ((double)g_attn_pin_bytes / total_bytes));
    bool whole = g_attn_part == 1 || (g_attn_part == 0 && nitems * 4 >= (long long)sms);
    // a capped grid that the items do not fill evenly would leave some CTAs a whole item more than others
    if (capped && g_attn_part == 0 && nitems > sms && nitems % sms != 0) whole = false;
    if (whole) {
        pt.align = spi;
        *grid = (int)(nitems < sms ? nitems : sms);
    } else {
        pt.align = 1;
        const int minshare = (spi + 7) / 8;
        long long g = pt.tot / minshare;
        if (g < 1) g = 1;
        if (g > sms) g = sms;
        *grid = (int)g;
    }
    return pt;
}

template <typename T>
inline bool fwd_eligible(const FwdParams& p) {
    if (!fwd_shape_ok(p.NB, p.L, p.D, p.H) || !p.s_part) return false;
    const size_t oa = sizeof(T) == 2 ? 8 : 16;
    return aligned(p.a, 16) && aligned(p.wa, 16) && aligned(p.uf, 16) && (p.ld_uf % 4 == 0) && aligned(p.v, 16) &&
           (!p.ctx || (aligned(p.ctx, oa) && p.ld_ctx % 4 == 0)) && (!p.z || (aligned(p.z, oa) && p.ld_z % 4 == 0)) &&
           (!p.ctx32 || (aligned(p.ctx32, 16) && p.ld_c32 % 4 == 0)) && aligned(p.s_part, 16);
}

template <typename T, int D, int H, int NBG, bool PRECISE>
int launch_fwd_inst(const FwdParams& p, cudaStream_t stream) {
    using G = Geo<T, D, H>;
    auto kern = fwd_stream_kernel<T, D, H, NBG, PRECISE>;
    static SatPerDevice configured;
    SAT_TRY(configured.once([&]() -> int {
        SAT_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_MAX));
        return SAT_OK;
    }));
    const Smem sl = fwd_smem<T, D, H, NBG>(p.L);
    SAT_REQUIRE(sl.nslots >= CWARPS && sl.total <= SMEM_MAX, "attention fwd (streaming): shared memory budget exceeded (L=%d)", p.L);
    const int ngroups = (p.NB + NBG - 1) / NBG, spi = (p.L + G::LS - 1) / G::LS;
    int grid = 1;
    const Part pt = pick_part((long long)p.NI * ngroups, spi, G::SLOT, ngroups, false, &grid, p.max_ctas);
    SatProfScope prof(SAT_PROF_ATTN_FWD, stream);
    cudaLaunchConfig_t cfg;
    cudaLaunchAttribute attr[1];
    pdl_config(&cfg, attr, dim3(grid), dim3(THREADS), sl.total, stream);
    SAT_CHECK_CUDA(cudaLaunchKernelEx(&cfg, kern, p, sl, pt, ngroups));
    if (pt.align == 1 && grid > 1) {
        fwd_combine_kernel<T, D, H, NBG, PRECISE><<<grid - 1, 256, 0, stream>>>(p, pt, grid, ngroups);
        SAT_CHECK_CUDA(cudaGetLastError());
    }
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

template <typename T>
inline bool bwd_eligible(const BwdParams& p) {
    if ((!p.dpre_out && !p.de_out) || p.da_acc || !p.ctx32 || !p.s_part || !shape_ok(p.L, p.D, p.H)) return false;
    if (!p.dz_is_dctx && !p.s_dctx) return false;
    const size_t oa = sizeof(T) == 2 ? 8 : 16;
    return aligned(p.a, 16) && aligned(p.wa, 16) && aligned(p.uf, 16) && (p.ld_uf % 4 == 0) && aligned(p.v, 16) &&
           aligned(p.dz, 16) && (p.ld_dz % 4 == 0) && (!p.dctx_extra || (aligned(p.dctx_extra, 16) && p.ld_dce % 4 == 0)) &&
           aligned(p.ctx32, 16) && (p.ld_c32 % 4 == 0) && aligned(p.duf, oa) && (p.ld_duf % 4 == 0) && aligned(p.dpre_out, 16) &&
           aligned(p.de_out, 4) && aligned(p.s_part, 16);
}

template <typename T, int D, int H, bool PRECISE>
int launch_bwd_inst(const BwdParams& p_in, cudaStream_t stream) {
    using G = Geo<T, D, H>;
    BwdParams p = p_in;
    SatProfScope prof(SAT_PROF_ATTN_BWD, stream);
    if (!p.dz_is_dctx) {
        dim3 g(sat_cdiv(D, 256), p.B);
        gate_bwd_kernel<T, PRECISE><<<g, 256, 0, stream>>>(p.dz, p.ld_dz, p.uf, p.ld_uf, p.dctx_extra, p.ld_dce, p.ctx32, p.ld_c32,
                                                          p.s_dctx, reinterpret_cast<T*>(p.duf), p.ld_duf, p.B, D, H, p.gate);
        SAT_CHECK_CUDA(cudaGetLastError());
        p.dz = p.s_dctx; p.ld_dz = D; p.dz_is_dctx = 1;
    }
    auto kern = bwd_stream_kernel<T, D, H, PRECISE>;
    static SatPerDevice configured;
    SAT_TRY(configured.once([&]() -> int {
        SAT_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_MAX));
        return SAT_OK;
    }));
    const Smem sl = bwd_smem<T, D, H>(p.L);
    SAT_REQUIRE(sl.nslots >= CWARPS && sl.total <= SMEM_MAX, "attention bwd (streaming): shared memory budget exceeded (L=%d)", p.L);
    const int spi = (p.L + G::LS - 1) / G::LS;
    int grid = 1;
    const Part pt = pick_part(p.B, spi, G::SLOT, 1, true, &grid, p.max_ctas);
    cudaLaunchConfig_t cfg;
    cudaLaunchAttribute attr[1];
    pdl_config(&cfg, attr, dim3(grid), dim3(THREADS), sl.total, stream);
    SAT_CHECK_CUDA(cudaLaunchKernelEx(&cfg, kern, p, sl, pt));
    if (pt.align == 1 && grid > 1) {
        bwd_combine_kernel<T, D, H><<<grid - 1, 256, 0, stream>>>(p, pt, grid);
        SAT_CHECK_CUDA(cudaGetLastError());
    }
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
    SAT_REQUIRE(p.ctx32 == nullptr, "attention fwd: ctx_f32 is only produced by the streaming kernel (shape/alignment/scratch do not qualify)");
    return attn::launch_fwd<T, PRECISE>(p, stream);
}
template <typename T, bool PRECISE>
int attn_step_bwd(const attn::BwdParams& p, cudaStream_t stream) {
    if (p.dpre_out || p.de_out) {
        SAT_REQUIRE(attn_s::bwd_eligible<T>(p), "attention bwd: dpre_out / de_out given but the streaming kernel does not support this shape/alignment/scratch");
        return attn_s::launch_bwd<T, PRECISE>(p, stream);
    }
    SAT_REQUIRE(p.dwa_acc != nullptr, "attention bwd: dwa_acc is NULL");
    return attn::launch_bwd<T, PRECISE>(p, stream);
}
