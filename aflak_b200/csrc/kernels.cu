// kernels.cu -- hand-written sm_100a kernels for aflak's data-parallel hot path.
//
// Every kernel here is HBM-bound streaming work (<= 2 flop/B): coalesced 16-byte lanes, enough independent
// loads in flight per thread to cover DRAM latency, grids sized from the SM count, no tensor cores.
// Rounding contract: every step the reference rounds to f32 is an explicit __f*_rn intrinsic here (the file
// is also compiled with -fmad=false), so elementwise results are bit-identical to the Rust code.
// Reference citations are relative to /root/reference/src/.
#include "kernels.h"

#include <cooperative_groups.h>

#include <atomic>
#include <cfloat>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <algorithm>

namespace afk {

static std::atomic<uint64_t> g_launches{0};
// Display kernels (tone curve, colour map, fused chain): 0 = fast forms that are proven / guarded to give the reference's
// indices and bytes (default), 1 = the reference's own IEEE sequence for every pixel (AFK_DISPLAY_EXACT=1, afk_set_display_exact)
static std::atomic<int> g_display_exact{-1};
bool display_exact() {
    int v = g_display_exact.load(std::memory_order_relaxed);
    if (v < 0) {
        const char *e = getenv("AFK_DISPLAY_EXACT");
        v = (e && atoi(e) != 0) ? 1 : 0;
        g_display_exact.store(v, std::memory_order_relaxed);
    }
    return v != 0;
}
void set_display_exact(bool on) { g_display_exact.store(on ? 1 : 0, std::memory_order_relaxed); }
uint64_t launch_count() { return g_launches.load(); }
#define AFK_COUNT_LAUNCH() g_launches.fetch_add(1, std::memory_order_relaxed)
void count_launch() { AFK_COUNT_LAUNCH(); }

// ----------------------------------------------------------------------------------------------------
// load helpers: streaming 16-byte loads that do not allocate in L1 (every voxel is touched once)
// ----------------------------------------------------------------------------------------------------
__device__ __forceinline__ float4 ld_stream(const float4 *p) {
    float4 r;
    asm("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
        : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
        : "l"(p));
    return r;
}
__device__ __forceinline__ float ld_stream(const float *p) {
    float r;
    asm("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(r) : "l"(p));
    return r;
}
__device__ __forceinline__ double2 ld_cg(const double2 *p) {
    double2 r;
    asm volatile("ld.global.cg.v2.f64 {%0,%1}, [%2];" : "=d"(r.x), "=d"(r.y) : "l"(p));
    return r;
}

template <int VEC> struct Lane;
template <> struct Lane<4> {
    using T = float4;
    static __device__ __forceinline__ void unpack(const float4 &v, float (&x)[4]) {
        x[0] = v.x; x[1] = v.y; x[2] = v.z; x[3] = v.w;
    }
    static __device__ __forceinline__ void store(float *p, const float (&x)[4]) {
        *reinterpret_cast<float4 *>(p) = make_float4(x[0], x[1], x[2], x[3]);
    }
};
template <> struct Lane<1> {
    using T = float;
    static __device__ __forceinline__ void unpack(const float &v, float (&x)[1]) { x[0] = v; }
    static __device__ __forceinline__ void store(float *p, const float (&x)[1]) { *p = x[0]; }
};
// store through an NVLink-SHARP multicast mapping: the switch replicates it into every bound GPU's memory
__device__ __forceinline__ void mc_store(float *p, const float (&x)[4]) {
    asm volatile("multimem.st.weak.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(x[0]), "f"(x[1]), "f"(x[2]), "f"(x[3]) : "memory");
}
__device__ __forceinline__ void mc_store(float *p, const float (&x)[1]) {
    asm volatile("multimem.st.weak.global.f32 [%0], %1;" ::"l"(p), "f"(x[0]) : "memory");
}

// ----------------------------------------------------------------------------------------------------
// K1/K2: axis-0 windowed reduction, stream-K decomposition
//
// Work is the flattened space (pixel tile p, concatenated window frame r): units = ntiles * ktot.  CTA g owns
// the contiguous unit range [g*q, (g+1)*q): every CTA streams the same number of tile-frames, so all SMs
// finish together whatever the cube shape (1024x1024x3840 or 74x74x4563).  A tile whose frames are owned by
// one CTA is finished in registers; a tile split between CTAs goes through f64 partials in scratch and the
// last CTA to arrive (per-tile counter) combines them in CTA order -> deterministic, no float atomics.
//
// Per pixel: compensated (Kahan) f32 accumulation in k order, plus the plain f32 sum `p` that carries
// NaN/Inf exactly like the reference's sequential sum (ndarray 0.12.1 sum_axis; lib.rs:1324-1330).
// ----------------------------------------------------------------------------------------------------
struct ReduceKArgs {
    const float *cube;
    int64_t n2;
    int nwin;
    int64_t start[kMaxWindows];
    int64_t woff[kMaxWindows + 1]; // prefix sum of window lengths; woff[nwin] == ktot
    float cnt[kMaxWindows];        // window length as the f32 the reference builds by repeated +1.0
    int64_t ktot, units, q, ntiles;
    int epilogue;
    float pa, pb, pfl, pmax, psign;
    float *out[kMaxPeers]; // nout destinations of the result (own buffer and/or peer-mapped buffers)
    int nout;
    unsigned *minmax;
    double2 *partials;
    unsigned *counters;
    // fused gather: when every CTA has stored its pixels, the last one writes `epoch` to each sig[i]
    unsigned *done;            // grid-wide arrival counter (self-resetting), or nullptr
    unsigned *sig[kMaxPeers];  // nsig flag words, typically one in every peer's flag array
    int nsig;
    unsigned epoch;
    int n_range_peer;
    unsigned *range_peer[kMaxPeers - 1];
    int mc; // out[0] is an NVLink multicast mapping: one multimem.st per tile instead of one store per GPU
    // cluster variant (small images): float4 lanes per pixel tile, frame sub-groups per CTA
    int tile4, jsub;
    // range memo: the image's {~key(min), key(max), any, 0} left beside the pixels, without atomics on the result and without
    // a zeroing pass: per-CTA slots + the last CTA to arrive (self-resetting counter) reduces them
    unsigned *range_out, *range_slots, *range_counter;
    unsigned long long *trace; // AFK_REDUCE_TRACE=1 (diagnostics): 4 globaltimer stamps per CTA, else nullptr
};

template <bool COMP> struct Acc {
    float s, c, p;
};

template <bool COMP> __device__ __forceinline__ void acc_init(Acc<COMP> &a) {
    // -0.0 is the additive identity for every x including -0.0 (0.0 + -0.0 would give +0.0)
    a.s = -0.0f;
    a.c = 0.0f;
    a.p = -0.0f;
}

template <bool COMP> __device__ __forceinline__ void acc_add(Acc<COMP> &a, float x) {
    a.p = __fadd_rn(a.p, x);
    if (COMP) {
        float y = __fsub_rn(x, a.c);
        float t = __fadd_rn(a.s, y);
        a.c = __fsub_rn(__fsub_rn(t, a.s), y);
        a.s = t;
    }
}

template <bool COMP> __device__ __forceinline__ double acc_value(const Acc<COMP> &a) {
    return COMP ? ((double)a.s - (double)a.c) : (double)a.p;
}

__device__ __forceinline__ float finalize_sum(double d, float p) {
    // non-finite plain sum: the reference's own f32 sum is NaN/Inf here -> reproduce it
    float r = (float)d;
    if (!(fabsf(p) <= FLT_MAX)) r = p;
    return r;
}

template <int VEC, bool COMP, int U>
__device__ __forceinline__ void acc_batch(const char *&ptr, size_t stride, Acc<COMP> (&acc)[VEC]) {
    using V = typename Lane<VEC>::T;
    V x[U];
#pragma unroll
    for (int j = 0; j < U; ++j) x[j] = ld_stream(reinterpret_cast<const V *>(ptr + (size_t)j * stride));
    ptr += (size_t)U * stride;
#pragma unroll
    for (int j = 0; j < U; ++j) {
        float e[VEC];
        Lane<VEC>::unpack(x[j], e);
#pragma unroll
        for (int i = 0; i < VEC; ++i) acc_add<COMP>(acc[i], e[i]);
    }
}

// frames [ka, kb) of one pixel lane, strictly in k order
template <int VEC, bool COMP, int U>
__device__ __forceinline__ void accumulate(const float *lane_base, int64_t n2, int64_t ka, int64_t kb,
                                           Acc<COMP> (&acc)[VEC]) {
    const size_t stride = (size_t)n2 * sizeof(float);
    const char *ptr = reinterpret_cast<const char *>(lane_base) + (size_t)ka * stride;
    int64_t n = kb - ka;
    for (; n >= U; n -= U) acc_batch<VEC, COMP, U>(ptr, stride, acc);
    if (U > 8 && n >= 8) { acc_batch<VEC, COMP, 8>(ptr, stride, acc); n -= 8; }
    if (U > 4 && n >= 4) { acc_batch<VEC, COMP, 4>(ptr, stride, acc); n -= 4; }
    if (n >= 2) { acc_batch<VEC, COMP, 2>(ptr, stride, acc); n -= 2; }
    if (n >= 1) { acc_batch<VEC, COMP, 1>(ptr, stride, acc); }
}

__device__ __forceinline__ unsigned f32_key(float f) {
    unsigned b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__host__ __device__ inline float key_f32(unsigned k) {
    unsigned b = (k & 0x80000000u) ? (k & 0x7FFFFFFFu) : ~k;
#ifdef __CUDA_ARCH__
    return __uint_as_float(b);
#else
    float f;
    memcpy(&f, &b, 4);
    return f;
#endif
}
float minmax_decode_min(unsigned inv_key) { return key_f32(~inv_key); }
float minmax_decode_max(unsigned key) { return key_f32(key); }

// Block-wide min/max of non-NaN values -> 3 atomics per CTA. result = {~key(min), key(max), any}
__device__ __forceinline__ void block_minmax_flush(float mn, float mx, bool any, unsigned *result) {
    __shared__ unsigned s_mn[32], s_mx[32], s_any[32];
    unsigned kmn = any ? ~f32_key(mn) : 0u, kmx = any ? f32_key(mx) : 0u, a = any ? 1u : 0u;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        kmn = max(kmn, __shfl_xor_sync(0xffffffffu, kmn, o));
        kmx = max(kmx, __shfl_xor_sync(0xffffffffu, kmx, o));
        a |= __shfl_xor_sync(0xffffffffu, a, o);
    }
    int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = (blockDim.x + 31) >> 5;
    if (l == 0) { s_mn[w] = kmn; s_mx[w] = kmx; s_any[w] = a; }
    __syncthreads();
    if (w == 0) {
        kmn = l < nw ? s_mn[l] : 0u;
        kmx = l < nw ? s_mx[l] : 0u;
        a = l < nw ? s_any[l] : 0u;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            kmn = max(kmn, __shfl_xor_sync(0xffffffffu, kmn, o));
            kmx = max(kmx, __shfl_xor_sync(0xffffffffu, kmx, o));
            a |= __shfl_xor_sync(0xffffffffu, a, o);
        }
        if (l == 0 && a) {
            atomicMax(result + 0, kmn);
            atomicMax(result + 1, kmx);
            atomicMax(result + 2, 1u);
        }
    }
}

// Range memo of a result image (image_min_max / get_vmin / get_vmax of what this launch wrote, NaN ignored): every CTA
// leaves its {~key(min), key(max), any} in its own slot; the last CTA to arrive reduces the slots and stores the image's four
// range words with plain stores.  Nothing has to be zeroed beforehand (the counter resets itself), so a consumer
// (convert_to_logscale of the template's tail, the display chain, image_min_max itself) finds the range on the device
// without a min/max pass or a memset.
__device__ __forceinline__ void block_range_reduce(unsigned &kmn, unsigned &kmx, unsigned &av, unsigned *s_mn, unsigned *s_mx, unsigned *s_any) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        kmn = max(kmn, __shfl_xor_sync(0xffffffffu, kmn, o));
        kmx = max(kmx, __shfl_xor_sync(0xffffffffu, kmx, o));
        av |= __shfl_xor_sync(0xffffffffu, av, o);
    }
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = (blockDim.x + 31) >> 5;
    if (l == 0) { s_mn[w] = kmn; s_mx[w] = kmx; s_any[w] = av; }
    __syncthreads();
    if (w == 0) {
        kmn = l < nw ? s_mn[l] : 0u;
        kmx = l < nw ? s_mx[l] : 0u;
        av = l < nw ? s_any[l] : 0u;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            kmn = max(kmn, __shfl_xor_sync(0xffffffffu, kmn, o));
            kmx = max(kmx, __shfl_xor_sync(0xffffffffu, kmx, o));
            av |= __shfl_xor_sync(0xffffffffu, av, o);
        }
    }
}
__device__ __forceinline__ void block_range_publish(float mn, float mx, bool any, unsigned *slots, unsigned *counter, unsigned *out) {
    __shared__ unsigned r_mn[32], r_mx[32], r_any[32];
    __shared__ int r_last;
    unsigned kmn = any ? ~f32_key(mn) : 0u, kmx = any ? f32_key(mx) : 0u, av = any ? 1u : 0u;
    block_range_reduce(kmn, kmx, av, r_mn, r_mx, r_any);
    if (threadIdx.x == 0) {
        unsigned *sl = slots + (size_t)blockIdx.x * 4;
        sl[0] = kmn; sl[1] = kmx; sl[2] = av;
        __threadfence();
        const unsigned prev = atomicAdd(counter, 1u);
        r_last = (prev == gridDim.x - 1);
    }
    __syncthreads();
    if (!r_last) return;
    __threadfence();
    kmn = 0u; kmx = 0u; av = 0u;
    for (unsigned i = threadIdx.x; i < gridDim.x; i += blockDim.x) {
        const volatile unsigned *sl = slots + (size_t)i * 4;
        kmn = max(kmn, sl[0]);
        kmx = max(kmx, sl[1]);
        av |= sl[2];
    }
    __syncthreads(); // r_mn.. are reused
    block_range_reduce(kmn, kmx, av, r_mn, r_mx, r_any);
    if (threadIdx.x == 0) {
        out[0] = kmn; out[1] = kmx; out[2] = av; out[3] = 0u;
        *counter = 0u; // self-reset for the next launch
    }
}

// create_equivalent_width, lib.rs:1584-1596: four separately rounded steps then the `> max` select
__device__ __forceinline__ float ew_value(float off, float on, float fl, float mx, float sign) {
    float t = __fsub_rn(off, on);
    t = __fmul_rn(t, fl);
    t = __fdiv_rn(t, off);
    t = __fmul_rn(t, sign);
    return (t > mx) ? 0.0f : t;
}

template <int NW>
__device__ __forceinline__ float reduce_epilogue(const ReduceKArgs &a, const float (&S)[NW]) {
    switch (a.epilogue) {
    case EPI_SUM: return S[0];
    case EPI_MEAN: return __fdiv_rn(S[0], a.cnt[0]);
    case EPI_MEAN_RATIO: {
        if (NW < 2) return 0.0f;
        float m0 = __fdiv_rn(S[0], a.cnt[0]);
        float m1 = __fdiv_rn(S[NW > 1 ? 1 : 0], a.cnt[1]);
        return __fdiv_rn(m0, m1); // &WcsArray / &WcsArray, unit.rs:863-880
    }
    case EPI_EW: {
        if (NW < 3) return 0.0f;
        float on = __fdiv_rn(S[0], a.cnt[0]);
        float off1 = __fdiv_rn(S[NW > 1 ? 1 : 0], a.cnt[1]);
        float off2 = __fdiv_rn(S[NW > 2 ? 2 : 0], a.cnt[2]);
        // linear_composition(u=off1, v=off2, a, b), lib.rs:1288-1297
        float off = __fadd_rn(__fmul_rn(off1, a.pa), __fmul_rn(off2, a.pb));
        return ew_value(off, on, a.pfl, a.pmax, a.psign);
    }
    }
    return 0.0f;
}

template <int VEC, bool COMP, int NW, int U, int MINB, int KG>
__global__ void __launch_bounds__(kReduceBlock, MINB) reduce_kernel(const ReduceKArgs a) {
    // KG k-groups of LANES threads share one pixel tile: group kg accumulates its own contiguous share of the segment's
    // frames, the groups are then combined through shared memory in group order (second level of the split-K; used for
    // small images, where there are far fewer pixel lanes than threads on the chip).  KG == 8: one warp per group.
    constexpr int LANES = kReduceBlock / KG;
    constexpr int64_t TILE = (int64_t)LANES * VEC;
    // a dependent elementwise kernel launched with programmatic stream serialization may be scheduled now; it blocks in
    // griddepcontrol.wait until this grid has completed and its memory is visible
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    const int tid = threadIdx.x;
    const int lane = tid % LANES, kg = tid / LANES;
    const int64_t g = blockIdx.x;
    int64_t u = g * a.q;
    const int64_t uend = min(u + a.q, a.units);
    const int64_t p_first = u / a.ktot;
    __shared__ int s_last;
    __shared__ double2 s_comb[KG > 1 ? KG : 1][KG > 1 ? LANES * VEC : 1];

    float mm_min = __int_as_float(0x7f800000), mm_max = __int_as_float(0xff800000);
    bool mm_any = false;

    while (u < uend) {
        const int64_t p = u / a.ktot;
        const int64_t tile_u0 = p * a.ktot;
        const int64_t r = u - tile_u0;
        const int64_t seg_end = min(uend, tile_u0 + a.ktot);
        const int64_t r_end = r + (seg_end - u);
        const bool full = (r == 0) && (r_end == a.ktot);
        const int64_t px = p * TILE + (int64_t)lane * VEC;
        const bool in_image = px < a.n2;
        const bool active = in_image && kg == 0; // the threads that own the tile's pixels after the group combine
        const float *lane_base = a.cube + px;
        const int slot = (p == p_first) ? 0 : 1;

        float S[NW][VEC];
#pragma unroll
        for (int w = 0; w < NW; ++w) {
            const int64_t lo = max(r, a.woff[w]), hi = min(r_end, a.woff[w + 1]);
            if (w < a.nwin && lo < hi) {
                Acc<COMP> acc[VEC];
#pragma unroll
                for (int i = 0; i < VEC; ++i) acc_init<COMP>(acc[i]);
                int64_t klo = lo, khi = hi;
                if (KG > 1) {
                    const int64_t c = (hi - lo + KG - 1) / KG;
                    klo = lo + (int64_t)kg * c;
                    khi = min(hi, klo + c);
                }
                if (in_image && klo < khi)
                    accumulate<VEC, COMP, U>(lane_base, a.n2, a.start[w] + (klo - a.woff[w]), a.start[w] + (khi - a.woff[w]), acc);
                double dv[VEC];
                float pv[VEC];
#pragma unroll
                for (int i = 0; i < VEC; ++i) { dv[i] = acc_value<COMP>(acc[i]); pv[i] = acc[i].p; }
                if (KG > 1) {
#pragma unroll
                    for (int i = 0; i < VEC; ++i) s_comb[kg][lane * VEC + i] = make_double2(dv[i], (double)pv[i]);
                    __syncthreads();
                    if (kg == 0) {
#pragma unroll
                        for (int i = 0; i < VEC; ++i) { dv[i] = -0.0; pv[i] = -0.0f; }
#pragma unroll 2
                        for (int j = 0; j < KG; ++j) {
#pragma unroll
                            for (int i = 0; i < VEC; ++i) {
                                const double2 v = s_comb[j][lane * VEC + i];
                                dv[i] += v.x;
                                pv[i] = __fadd_rn(pv[i], (float)v.y);
                            }
                        }
                    }
                    __syncthreads(); // s_comb is reused by the next window
                }
                if (full) {
#pragma unroll
                    for (int i = 0; i < VEC; ++i) S[w][i] = finalize_sum(dv[i], pv[i]);
                } else if (active) {
                    double2 *dst = a.partials + (((size_t)g * 2 + slot) * a.nwin + w) * TILE + (size_t)lane * VEC;
#pragma unroll
                    for (int i = 0; i < VEC; ++i) dst[i] = make_double2(dv[i], (double)pv[i]);
                }
            }
        }

        if (!full) {
            // publish partials, then count arrivals on this tile; the last CTA to arrive combines
            __threadfence();
            __syncthreads();
            if (tid == 0) {
                const int64_t g_lo = tile_u0 / a.q, g_hi = (tile_u0 + a.ktot - 1) / a.q;
                const unsigned expect = (unsigned)(g_hi - g_lo + 1);
                const unsigned prev = atomicAdd(a.counters + p, 1u);
                const int last = (prev == expect - 1);
                if (last) a.counters[p] = 0u; // self-reset for the next launch
                s_last = last;
            }
            __syncthreads();
            const bool last = s_last != 0;
            if (!last) {
                u = seg_end;
                continue;
            }
            __threadfence();
            const int64_t g_lo = tile_u0 / a.q, g_hi = (tile_u0 + a.ktot - 1) / a.q;
#pragma unroll
            for (int w = 0; w < NW; ++w) {
                if (w < a.nwin) {
                    double d[VEC];
                    float pl[VEC];
#pragma unroll
                    for (int i = 0; i < VEC; ++i) { d[i] = -0.0; pl[i] = -0.0f; }
                    // group kg takes the contributing CTAs g_lo+kg, g_lo+kg+KG, ... (all of them when KG == 1)
                    for (int64_t gg = g_lo + kg; gg <= g_hi; gg += KG) {
                        const int64_t cs = max(gg * a.q, tile_u0) - tile_u0;
                        const int64_t ce = min((gg + 1) * a.q, tile_u0 + a.ktot) - tile_u0;
                        if (max(cs, a.woff[w]) < min(ce, a.woff[w + 1])) {
                            const int sl = ((gg * a.q) / a.ktot == p) ? 0 : 1;
                            if (in_image) {
                                const double2 *src = a.partials + (((size_t)gg * 2 + sl) * a.nwin + w) * TILE + (size_t)lane * VEC;
                                double2 v[VEC];
#pragma unroll
                                for (int i = 0; i < VEC; ++i) v[i] = ld_cg(src + i);
#pragma unroll
                                for (int i = 0; i < VEC; ++i) {
                                    d[i] += v[i].x;
                                    pl[i] = __fadd_rn(pl[i], (float)v[i].y);
                                }
                            }
                        }
                    }
                    if (KG > 1) {
#pragma unroll
                        for (int i = 0; i < VEC; ++i) s_comb[kg][lane * VEC + i] = make_double2(d[i], (double)pl[i]);
                        __syncthreads();
                        if (kg == 0) {
#pragma unroll
                            for (int i = 0; i < VEC; ++i) { d[i] = -0.0; pl[i] = -0.0f; }
#pragma unroll 2
                            for (int j = 0; j < KG; ++j) {
#pragma unroll
                                for (int i = 0; i < VEC; ++i) {
                                    const double2 v = s_comb[j][lane * VEC + i];
                                    d[i] += v.x;
                                    pl[i] = __fadd_rn(pl[i], (float)v.y);
                                }
                            }
                        }
                        __syncthreads();
                    }
#pragma unroll
                    for (int i = 0; i < VEC; ++i) S[w][i] = finalize_sum(d[i], pl[i]);
                }
            }
        }

        if (active) {
            float o[VEC];
#pragma unroll
            for (int i = 0; i < VEC; ++i) {
                float t[NW];
#pragma unroll
                for (int w = 0; w < NW; ++w) t[w] = S[w][i];
                o[i] = reduce_epilogue<NW>(a, t);
                if (o[i] == o[i]) { // Rust f32::min/max ignore a NaN operand (lib.rs:186-195)
                    mm_min = fminf(mm_min, o[i]);
                    mm_max = fmaxf(mm_max, o[i]);
                    mm_any = true;
                }
            }
            if (a.mc) {
                mc_store(a.out[0] + px, o);
            } else {
#pragma unroll 1
                for (int d = 0; d < a.nout; ++d) Lane<VEC>::store(a.out[d] + px, o);
            }
        }
        u = seg_end;
    }
    if (a.minmax) block_minmax_flush(mm_min, mm_max, mm_any, a.minmax);
    if (a.range_out) block_range_publish(mm_min, mm_max, mm_any, a.range_slots, a.range_counter, a.range_out);
    if (a.done) {
        // Fused gather epilogue: make this CTA's (peer) stores visible system-wide, count the arrival; the last CTA
        // of the grid publishes the epoch to every destination's flag word with release semantics.
        __threadfence_system();
        __syncthreads();
        if (tid == 0) {
            const unsigned prev = atomicAdd(a.done, 1u);
            if (prev == gridDim.x - 1) {
                *a.done = 0u; // self-reset for the next launch
                if (a.range_out) {
                    // min/max exchange: every CTA has been through block_range_publish, so the band's range words are final
                    __threadfence();
                    unsigned rw[4];
#pragma unroll
                    for (int i = 0; i < 4; ++i) rw[i] = *reinterpret_cast<volatile unsigned *>(a.range_out + i);
                    for (int d = 0; d < a.n_range_peer; ++d)
                        *reinterpret_cast<uint4 *>(a.range_peer[d]) = make_uint4(rw[0], rw[1], rw[2], rw[3]);
                }
                __threadfence_system();
                for (int d = 0; d < a.nsig; ++d)
                    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(a.sig[d]), "r"(a.epoch) : "memory");
            }
        }
    }
}

// ----------------------------------------------------------------------------------------------------
// Small images (MaNGA 74x74x4563: far fewer pixel lanes than threads on the chip): thread-block clusters.
//
// One cluster owns one pixel tile of `tile4` float4 lanes at a time; the tile's concatenated window frames are
// split evenly over the cluster's CTAs, and inside a CTA over `jsub` interleaved frame sub-groups (thread =
// (lane, sub-group), so a warp's 16-byte loads of one frame are contiguous).  Every thread keeps U independent
// loads in flight; the remainder is one predicated batch (one memory latency, not log2(U) of them).
// Merge: sub-groups through shared memory, then the CTAs of the cluster through DISTRIBUTED shared memory --
// CTA r finishes pixel slice r of the tile from every CTA's partials in rank order (deterministic).  No global
// scratch, no arrival counters, no fences: this replaces the f64-partials round trip of reduce_kernel, which
// costs ~8 us of a 33 us launch at this size (tools/c1_probe.cu, profiles/r2_c1_probe.txt).
// ----------------------------------------------------------------------------------------------------
namespace cg = cooperative_groups;

__device__ __forceinline__ void cl_stamp(const ReduceKArgs &a, int slot) {
    if (a.trace && threadIdx.x == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        a.trace[(size_t)blockIdx.x * 4 + slot] = t;
    }
}
template <int H> __device__ __forceinline__ void cl_load(float4 (&x)[H], const char *ptr, size_t stride, int64_t n) {
#pragma unroll
    for (int u = 0; u < H; ++u)
        if (u < n) x[u] = ld_stream(reinterpret_cast<const float4 *>(ptr + (size_t)u * stride));
}
template <int H> __device__ __forceinline__ void cl_add(Acc<true> (&acc)[4], const float4 (&x)[H], int64_t n) {
#pragma unroll
    for (int u = 0; u < H; ++u)
        if (u < n) {
            acc_add<true>(acc[0], x[u].x);
            acc_add<true>(acc[1], x[u].y);
            acc_add<true>(acc[2], x[u].z);
            acc_add<true>(acc[3], x[u].w);
        }
}
// the same pixel of every CTA's partial, all remote loads issued before the first add (rank order kept in the sum)
template <int CS> __device__ __forceinline__ void cl_gather(cg::cluster_group &cl, double2 *part, size_t idx, int C, double &d, float &pl) {
    for (int c0 = 0; c0 < C; c0 += CS) {
        double2 v[CS];
#pragma unroll
        for (int c = 0; c < CS; ++c)
            if (c0 + c < C) v[c] = cl.map_shared_rank(part, c0 + c)[idx];
#pragma unroll
        for (int c = 0; c < CS; ++c)
            if (c0 + c < C) {
                d += v[c].x;
                pl = __fadd_rn(pl, (float)v[c].y);
            }
    }
}

template <int NW, int U>
__global__ void __launch_bounds__(kReduceBlock, 2) reduce_cluster_kernel(const ReduceKArgs a) {
    extern __shared__ __align__(16) unsigned char cl_smem[];
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    cg::cluster_group cl = cg::this_cluster();
    const int C = (int)cl.num_blocks(), rank = (int)cl.block_rank();
    const int ncl = (int)gridDim.x / C, cid = (int)blockIdx.x / C;
    const int tid = threadIdx.x;
    const int li = tid % a.tile4, j = tid / a.tile4;
    const int tpx = a.tile4 * 4;
    double2 *part = reinterpret_cast<double2 *>(cl_smem); // [NW][tpx]   this CTA's share of the tile, read by the cluster
    double2 *comb = part + (size_t)NW * tpx;              // [jsub][NW][tpx]
    const int64_t kc = (a.ktot + C - 1) / C;
    const int64_t r0 = (int64_t)rank * kc, r1 = min(a.ktot, r0 + kc);
    const int64_t n4tot = a.n2 / 4;

    float mm_min = __int_as_float(0x7f800000), mm_max = __int_as_float(0xff800000);
    bool mm_any = false;
    cl_stamp(a, 0);

    for (int64_t tile = cid; tile < a.ntiles; tile += ncl) {
        const int64_t l0 = tile * a.tile4;
        const int n4 = (int)min((int64_t)a.tile4, n4tot - l0);
        const bool active = (j < a.jsub) && (li < n4);
        // every thread leaves its (f64 value, plain f32 sum) per window in shared memory right away: nothing is carried in
        // registers from one window to the next
        double2 *mine = (a.jsub > 1 ? comb + (size_t)j * NW * tpx : part) + li * 4;
#pragma unroll
        for (int w = 0; w < NW; ++w) {
            double dv[4];
            float pv[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) { dv[e] = -0.0; pv[e] = -0.0f; }
            const int64_t lo = max(r0, a.woff[w]) + j, hi = min(r1, a.woff[w + 1]);
            if (active && lo < hi) {
                Acc<true> acc[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) acc_init<true>(acc[e]);
                int64_t n = (hi - lo + a.jsub - 1) / a.jsub; // this thread's frames: lo, lo + jsub, ...
                const size_t stride = (size_t)a.n2 * sizeof(float) * (size_t)a.jsub;
                const char *ptr = reinterpret_cast<const char *>(a.cube + (a.start[w] + (lo - a.woff[w])) * a.n2 + (l0 + li) * 4);
                // two half-batches in ping-pong: the loads of the next half are in flight while this half is added
                constexpr int H = U / 2;
                float4 xa[H], xb[H];
                cl_load<H>(xa, ptr, stride, n);
                while (n > 0) {
                    cl_load<H>(xb, ptr + (size_t)H * stride, stride, n - H);
                    cl_add<H>(acc, xa, n);
                    if (n <= H) break;
                    cl_load<H>(xa, ptr + (size_t)U * stride, stride, n - U);
                    cl_add<H>(acc, xb, n - H);
                    n -= U;
                    ptr += (size_t)U * stride;
                }
#pragma unroll
                for (int e = 0; e < 4; ++e) { dv[e] = acc_value<true>(acc[e]); pv[e] = acc[e].p; }
            }
            if (active) {
#pragma unroll
                for (int e = 0; e < 4; ++e) mine[(size_t)w * tpx + e] = make_double2(dv[e], (double)pv[e]);
            }
        }
        // sub-groups -> this CTA's partial, in sub-group order
        if (a.jsub > 1) {
            __syncthreads();
            cl_stamp(a, 1);
            if (active && j == 0) {
#pragma unroll
                for (int w = 0; w < NW; ++w)
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        double d = -0.0;
                        float pl = -0.0f;
                        for (int jj = 0; jj < a.jsub; ++jj) {
                            const double2 v = comb[((size_t)jj * NW + w) * tpx + li * 4 + e];
                            d += v.x;
                            pl = __fadd_rn(pl, (float)v.y);
                        }
                        part[(size_t)w * tpx + li * 4 + e] = make_double2(d, (double)pl);
                    }
            }
        }
        cl.sync();
        cl_stamp(a, 2);
        // CTA `rank` finishes pixel slice `rank` of the tile from every CTA's partial, in rank order.  tile4 <= 64 lanes
        // (reduce_plan_cluster), so a slice is at most 256 / C pixels: one pixel per thread.
        {
            const int npx = n4 * 4;
            const int pc = (npx + C - 1) / C;
            const int q = rank * pc + tid;
            const bool mine_px = tid < pc && q < npx;
            float S[NW];
            if (mine_px) {
#pragma unroll
                for (int w = 0; w < NW; ++w) {
                    double d = -0.0;
                    float pl = -0.0f;
                    cl_gather<4>(cl, part, (size_t)w * tpx + q, C, d, pl);
                    S[w] = finalize_sum(d, pl);
                }
            }
            // the remote partials are in registers: let the other CTAs run ahead (their partials may be overwritten by the
            // next tile, or their CTA may exit, only after every reader has arrived)
            asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
            if (mine_px) {
                const float o = reduce_epilogue<NW>(a, S);
                if (o == o) {
                    mm_min = fminf(mm_min, o);
                    mm_max = fmaxf(mm_max, o);
                    mm_any = true;
                }
                a.out[0][l0 * 4 + q] = o;
            }
            asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
        }
        cl_stamp(a, 3);
    }
    if (a.minmax) block_minmax_flush(mm_min, mm_max, mm_any, a.minmax);
    if (a.range_out) block_range_publish(mm_min, mm_max, mm_any, a.range_slots, a.range_counter, a.range_out);
}

// Stream-ordered wait of the fused gather: lane i spins until flags[i] >= epoch (epochs are monotonic), with a
// clock bound so a lost peer turns into an error word and a NaN band instead of a hung GPU or silently stale pixels.
__global__ void gather_wait_kernel(const GatherWaitArgs a, long long max_cycles) {
    const int i = threadIdx.x;
    bool timed_out = false;
    if (i < a.n) {
        const long long t0 = clock64();
        for (;;) {
            unsigned v;
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(a.flags + i) : "memory");
            if ((int)(v - a.epoch) >= 0) break;
            if (clock64() - t0 > max_cycles) {
                atomicMax(a.err, 1u + (unsigned)i);
                timed_out = true;
                break;
            }
            __nanosleep(200);
        }
    }
    unsigned bad = __ballot_sync(0xffffffffu, timed_out);
    if (a.combined) {
        // min/max exchange: the bands' ranges arrived before their flags (release / acquire); combine them in the warp
        unsigned kmn = 0u, kmx = 0u, av = 0u;
        if (i < a.n && !timed_out) {
            const volatile unsigned *rp = a.ranges + 4 * i;
            const uint3 r = make_uint3(rp[0], rp[1], rp[2]);
            kmn = r.x; kmx = r.y; av = r.z;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            kmn = max(kmn, __shfl_xor_sync(0xffffffffu, kmn, o));
            kmx = max(kmx, __shfl_xor_sync(0xffffffffu, kmx, o));
            av |= __shfl_xor_sync(0xffffffffu, av, o);
        }
        if (i == 0) *reinterpret_cast<uint4 *>(a.combined) = make_uint4(kmn, kmx, av, 0u);
    }
    while (bad) { // error path only: the whole warp poisons the band that never arrived
        const int r = __ffs(bad) - 1;
        bad &= bad - 1;
        float *p = a.image + a.band_off[r];
        for (int64_t j = i; j < a.band_len[r]; j += 32) p[j] = __int_as_float(0x7fc00000);
    }
}

// Band broadcast: 16-byte words of this rank's band to every peer's copy, then the epoch flags (same arrival protocol as
// the fused gather of reduce_kernel).
__global__ void __launch_bounds__(256) band_broadcast_kernel(const BandBroadcastArgs a) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < a.n16; i += stride) {
        const uint4 v = a.src[i];
        if (a.mc_dst) { // one store through the multicast mapping (bits are copied, not interpreted)
            asm volatile("multimem.st.weak.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(a.mc_dst + i), "f"(__uint_as_float(v.x)),
                         "f"(__uint_as_float(v.y)), "f"(__uint_as_float(v.z)), "f"(__uint_as_float(v.w))
                         : "memory");
        } else {
            for (int d = 0; d < a.ndst; ++d) a.dst[d][i] = v;
        }
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned prev = atomicAdd(a.done, 1u);
        if (prev == gridDim.x - 1) {
            *a.done = 0u;
            __threadfence_system();
            for (int d = 0; d < a.nsig; ++d)
                asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(a.sig[d]), "r"(a.epoch) : "memory");
        }
    }
}

cudaError_t band_broadcast_launch(const BandBroadcastArgs &a, int sm_count, cudaStream_t s) {
    if (a.ndst < 0 || a.ndst > kMaxPeers - 1 || a.nsig < 1 || a.nsig > kMaxPeers || !a.done) return cudaErrorInvalidValue;
    int64_t need = (a.n16 + 255) / 256;
    if (need < 1) need = 1;
    const int64_t cap = (int64_t)sm_count * 4;
    band_broadcast_kernel<<<(int)(need < cap ? need : cap), 256, 0, s>>>(a);
    AFK_COUNT_LAUNCH();
    return cudaGetLastError();
}

cudaError_t gather_wait_launch(const GatherWaitArgs &a, double timeout_s, int device, cudaStream_t s) {
    if (a.n < 1 || a.n > kMaxPeers) return cudaErrorInvalidValue;
    // cudaDevAttrClockRate is not cached by the runtime (~1 ms per query): ask once per device
    static std::atomic<int> khz_of[16];
    int khz = (device >= 0 && device < 16) ? khz_of[device].load(std::memory_order_relaxed) : 0;
    if (khz <= 0) {
        if (cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, device) != cudaSuccess || khz <= 0) khz = 1965000;
        if (device >= 0 && device < 16) khz_of[device].store(khz, std::memory_order_relaxed);
    }
    const long long cycles = (long long)(timeout_s * 1e3 * (double)khz);
    gather_wait_kernel<<<1, 32, 0, s>>>(a, cycles);
    AFK_COUNT_LAUNCH();
    return cudaGetLastError();
}

static float count_as_f32(int64_t n) {
    // ndarray 0.12.1 mean_axis builds the divisor by n-1 additions of 1.0f32; exact up to 2^24
    if (n <= (1 << 24)) return (float)n;
    float c = 1.0f;
    for (int64_t i = 1; i < n; ++i) c = c + 1.0f;
    return c;
}

int reduce_tune_ctas();
int reduce_tune_kg();
static int env_int(const char *name, int dflt);

constexpr int kClusterMaxTile4 = 128;                 // float4 lanes per tile (<= kReduceBlock)
constexpr size_t kClusterMaxSmem = (size_t)(kMaxWindows * kClusterMaxTile4 * 4 + kMaxWindows * kReduceBlock * 4) * sizeof(double2);

template <typename K> static cudaError_t cluster_kernel_attrs(K kernel) {
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kClusterMaxSmem);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    return e;
}

// How many clusters of `csize` CTAs the device keeps resident at once (worst-case shared memory); cached per size.
static int cluster_capacity(int csize) {
    static std::atomic<int> cache[17];
    if (csize < 1 || csize > 16) return 0;
    int v = cache[csize].load(std::memory_order_relaxed);
    if (v != 0) return v < 0 ? 0 : v;
    int n = 0;
    if (cluster_kernel_attrs(reduce_cluster_kernel<1, 16>) == cudaSuccess) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(csize * 64);
        cfg.blockDim = dim3(kReduceBlock);
        cfg.dynamicSmemBytes = kClusterMaxSmem;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = csize;
        at[0].val.clusterDim.y = 1;
        at[0].val.clusterDim.z = 1;
        cfg.attrs = at;
        cfg.numAttrs = 1;
        if (cudaOccupancyMaxActiveClusters(&n, reduce_cluster_kernel<1, 16>, &cfg) != cudaSuccess) n = 0;
    }
    cudaGetLastError();
    cache[csize].store(n > 0 ? n : -1, std::memory_order_relaxed);
    return n;
}

// AFK_REDUCE_CLUSTER: 0 disables the cluster variant, N = 2|4|8|16 sets the cluster size (default 8).
static void reduce_plan_cluster(const ReduceArgs &a, ReducePlan &pl) {
    static const int want = env_int("AFK_REDUCE_CLUSTER", 8);
    static const int jmax = env_int("AFK_REDUCE_CLUSTER_JMAX", 8);
    if (want < 1 || a.mode == SUM_SEQUENTIAL) return;
    // template-sized windows (3 x 20 frames, 1.3 MB) are one load batch either way: the cluster launch and its two
    // barriers cost more than they save (measured 16.4 vs 14.3 us, profiles/r2_cluster_tuning.jsonl)
    if (pl.ktot < 128) return;
    int C = want > 16 ? 16 : want;
    const int64_t n4tot = a.n2 / 4;
    int cap = 0;
    int64_t tile4 = 0;
    int jsub = 1;
    for (;; C /= 2) {
        cap = cluster_capacity(C);
        if (cap > 0) {
            const int64_t waves = (n4tot + (int64_t)cap * kClusterMaxTile4 - 1) / ((int64_t)cap * kClusterMaxTile4);
            tile4 = (n4tot + cap * waves - 1) / (cap * waves);
            if (tile4 < 1) tile4 = 1;
            jsub = (int)(kReduceBlock / tile4);
            if (jsub > jmax) jsub = jmax;
            if (jsub < 1) jsub = 1;
            // every (CTA, sub-group) should own at least two frames; otherwise a smaller cluster does the same work
            if (C == 1 || pl.ktot >= (int64_t)C * jsub * 2) break;
        }
        if (C == 1) return; // no cluster launch possible on this device/context: keep the stream-K kernel
    }
    // Measured (profiles/r2_cluster_tuning.jsonl): the variant wins while a tile is narrow enough for >= 4 frame sub-groups
    // per CTA (74x74: 25.0 vs 28.1 us back to back); at 128x128 it ties and at 300x300 the stream-K kernel is 20 % faster.
    if (C < 4 || tile4 > kReduceBlock / 4) return;
    const int64_t ntiles = (n4tot + tile4 - 1) / tile4;
    const int64_t ncl = ntiles < cap ? ntiles : cap;
    pl.cluster = C;
    pl.tile4 = (int)tile4;
    pl.jsub = jsub;
    pl.cl_unroll = 16;
    pl.ntiles = ntiles;
    pl.units = ntiles * pl.ktot;
    pl.q = (pl.ktot + C - 1) / C;
    pl.grid = (int)(ncl * C);
    pl.cl_smem = (size_t)(a.nwin * tile4 * 4 + (jsub > 1 ? (size_t)jsub * a.nwin * tile4 * 4 : 0)) * sizeof(double2);
    pl.partial_bytes = 0;
    pl.n_counters = 0;
}

ReducePlan reduce_plan(const ReduceArgs &a, int sm_count) {
    ReducePlan pl{};
    bool aligned = (a.n2 % 4 == 0) && ((reinterpret_cast<uintptr_t>(a.cube) & 15u) == 0) &&
                   ((reinterpret_cast<uintptr_t>(a.out) & 15u) == 0);
    for (int d = 0; d < a.n_extra_out; ++d) aligned = aligned && ((reinterpret_cast<uintptr_t>(a.extra_out[d]) & 15u) == 0);
    pl.vec = aligned ? 4 : 1;
    int64_t G = (int64_t)sm_count * reduce_tune_ctas();
    // Small images (fewer pixel tiles than half the CTAs of a full wave, e.g. MaNGA 74x74): split the spectral axis
    // inside the CTA as well -- 8 k-groups of one warp each per 32-lane tile.  Never in sequential mode.
    const int64_t tiles_kg1 = (a.n2 + (int64_t)kReduceBlock * pl.vec - 1) / ((int64_t)kReduceBlock * pl.vec);
    pl.kg = (a.mode != SUM_SEQUENTIAL && tiles_kg1 * 2 < G && reduce_tune_kg()) ? kReduceKGroups : 1;
    if (pl.kg > 1) G = (int64_t)sm_count * kReduceKgCtasPerSm; // that variant is compiled for 2 resident CTAs per SM
    const int64_t tile = (int64_t)(kReduceBlock / pl.kg) * pl.vec;
    pl.ntiles = (a.n2 + tile - 1) / tile;
    pl.ktot = 0;
    for (int w = 0; w < a.nwin; ++w) pl.ktot += a.end[w] - a.start[w];
    pl.units = pl.ntiles * pl.ktot;
    if (a.mode == SUM_SEQUENTIAL) {
        pl.q = pl.ktot * ((pl.ntiles + G - 1) / G); // whole tiles only: never split a pixel's sum
    } else {
        int64_t q = (pl.units + G - 1) / G;
        const int64_t want = pl.kg > 1 ? 4 * pl.kg : 8; // at least a few frames per k-group / one 8-frame batch per CTA
        const int64_t qmin = pl.ktot < want ? pl.ktot : want;
        pl.q = q < qmin ? qmin : q;
        // a tile with little work (template-sized windows on a small image) is not worth a cross-CTA split: the
        // partials round trip and the arrival counter cost more than the frames
        if (pl.kg > 1 && pl.q < pl.ktot && pl.ktot <= 32 * pl.kg) pl.q = pl.ktot;
    }
    if (pl.q < 1) pl.q = 1;
    pl.grid = (int)((pl.units + pl.q - 1) / pl.q);
    pl.partial_bytes = (size_t)pl.grid * 2 * a.nwin * tile * sizeof(double2);
    pl.n_counters = (size_t)pl.ntiles;
    if (pl.kg > 1 && pl.vec == 4 && a.n_signal == 0 && a.n_extra_out == 0 && !a.multicast_out) reduce_plan_cluster(a, pl);
    // range memo: 4 words per CTA after the partials, one more self-resetting counter after the tile counters
    pl.range_slot_offset = (pl.partial_bytes + 15) / 16 * 16;
    pl.partial_bytes = pl.range_slot_offset + (size_t)pl.grid * 4 * sizeof(unsigned);
    pl.range_counter_index = pl.n_counters;
    pl.n_counters += 1;
    return pl;
}

// Tuning knobs, read once (AFK_REDUCE_UNROLL = 8|16 independent 16-byte loads per thread, AFK_REDUCE_CTAS = 2|3|4
// resident CTAs per SM).  Measured on B200, ALMA-shape cube (profiles/r1_reduce_tuning_sweep.txt): (8,3) 7.10 TB/s,
// (16,3) 7.08, (8,2) 7.01, (16,2) 7.00, (8,4) 6.72 (64 registers: spills + serialised loads), (4,4) 5.62.
static int env_int(const char *name, int dflt) {
    const char *v = getenv(name);
    return v ? atoi(v) : dflt;
}
int reduce_tune_unroll() { static int u = env_int("AFK_REDUCE_UNROLL", 8); return u == 16 ? 16 : 8; }
int reduce_tune_ctas() { static int c = env_int("AFK_REDUCE_CTAS", kReduceCtasPerSm); return (c == 2 || c == 4) ? c : 3; }
int reduce_tune_kg() { static int k = env_int("AFK_REDUCE_KGROUPS", 1); return k != 0; } // 0 disables the in-CTA k split

template <int VEC, bool COMP, int U, int MINB, int KG = 1>
static cudaError_t reduce_dispatch_nw(const ReduceKArgs &k, int grid, cudaStream_t s) {
    switch (k.nwin) {
    case 1: reduce_kernel<VEC, COMP, 1, U, MINB, KG><<<grid, kReduceBlock, 0, s>>>(k); break;
    case 2: reduce_kernel<VEC, COMP, 2, U, MINB, KG><<<grid, kReduceBlock, 0, s>>>(k); break;
    case 3: reduce_kernel<VEC, COMP, 3, U, MINB, KG><<<grid, kReduceBlock, 0, s>>>(k); break;
    default: return cudaErrorInvalidValue;
    }
    AFK_COUNT_LAUNCH();
    return cudaGetLastError();
}

template <int VEC, bool COMP>
static cudaError_t reduce_dispatch_tune(const ReduceKArgs &k, int grid, int kg, cudaStream_t s) {
    if (kg > 1) { // small images are latency-bound: AFK_REDUCE_KG_UNROLL=16 doubles the loads a warp keeps in flight
        static const int ukg = env_int("AFK_REDUCE_KG_UNROLL", 8);
        if (ukg == 16) return reduce_dispatch_nw<VEC, COMP, 16, kReduceKgCtasPerSm, kReduceKGroups>(k, grid, s);
        return reduce_dispatch_nw<VEC, COMP, 8, kReduceKgCtasPerSm, kReduceKGroups>(k, grid, s);
    }
    const int u = reduce_tune_unroll(), c = reduce_tune_ctas();
#define AFK_TUNE(U_, C_) if (u == U_ && c == C_) return reduce_dispatch_nw<VEC, COMP, U_, C_>(k, grid, s)
    AFK_TUNE(8, 4); AFK_TUNE(8, 2); AFK_TUNE(16, 2); AFK_TUNE(16, 3);
#undef AFK_TUNE
    return reduce_dispatch_nw<VEC, COMP, 8, 3>(k, grid, s);
}

// Diagnostics (AFK_REDUCE_TRACE=1): per-CTA globaltimer stamps -> phase durations on stderr.  Synchronises.
template <typename K> static cudaError_t reduce_cluster_traced(cudaLaunchConfig_t &cfg, K kernel, ReduceKArgs k) {
    const size_t n = (size_t)cfg.gridDim.x * 4;
    unsigned long long *d = nullptr;
    cudaError_t e = cudaMalloc(&d, n * 8);
    if (e != cudaSuccess) return e;
    cudaMemset(d, 0, n * 8);
    k.trace = d;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaEventRecord(e0, cfg.stream);
    e = cudaLaunchKernelEx(&cfg, kernel, k);
    cudaEventRecord(e1, cfg.stream);
    AFK_COUNT_LAUNCH();
    cudaStreamSynchronize(cfg.stream);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    std::vector<unsigned long long> h(n);
    cudaMemcpy(h.data(), d, n * 8, cudaMemcpyDeviceToHost);
    cudaFree(d);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    unsigned long long lo[4], hi[4];
    for (int sl = 0; sl < 4; ++sl) { lo[sl] = ~0ull; hi[sl] = 0; }
    for (size_t i = 0; i < n; ++i) {
        if (!h[i]) continue;
        lo[i & 3] = std::min(lo[i & 3], h[i]);
        hi[i & 3] = std::max(hi[i & 3], h[i]);
    }
    const double t0 = (double)lo[0];
    fprintf(stderr, "[afk trace] grid %u events %.2f us | start %.2f..%.2f | stream done %.2f..%.2f | cluster synced %.2f..%.2f | end %.2f..%.2f (us after the first CTA start)\n",
            cfg.gridDim.x, ms * 1e3, 0.0, (hi[0] - t0) / 1e3, (lo[1] - t0) / 1e3, (hi[1] - t0) / 1e3, (lo[2] - t0) / 1e3, (hi[2] - t0) / 1e3,
            (lo[3] - t0) / 1e3, (hi[3] - t0) / 1e3);
    return e;
}

template <int NW> static cudaError_t reduce_cluster_go(const ReduceKArgs &k, const ReducePlan &pl, cudaStream_t s) {
    static const cudaError_t attrs = cluster_kernel_attrs(reduce_cluster_kernel<NW, 16>); // once per instantiation
    if (attrs != cudaSuccess) return attrs;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)pl.grid);
    cfg.blockDim = dim3(kReduceBlock);
    cfg.dynamicSmemBytes = pl.cl_smem;
    cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = (unsigned)pl.cluster;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    static const int trace = env_int("AFK_REDUCE_TRACE", 0);
    if (trace) return reduce_cluster_traced(cfg, reduce_cluster_kernel<NW, 16>, k);
    const cudaError_t e = cudaLaunchKernelEx(&cfg, reduce_cluster_kernel<NW, 16>, k);
    AFK_COUNT_LAUNCH();
    return e;
}

static cudaError_t reduce_cluster_dispatch(const ReduceKArgs &k, const ReducePlan &pl, cudaStream_t s) {
    switch (k.nwin) {
    case 1: return reduce_cluster_go<1>(k, pl, s);
    case 2: return reduce_cluster_go<2>(k, pl, s);
    case 3: return reduce_cluster_go<3>(k, pl, s);
    }
    return cudaErrorInvalidValue;
}

void reduce_describe(const ReduceArgs &a, const ReducePlan &pl, char *buf, size_t cap) {
    if (pl.cluster > 0) {
        snprintf(buf, cap, "reduce_cluster_kernel<NW=%d,U=%d> grid=%d block=%d cluster=%d tiles=%lld lanes_per_tile=%d frame_subgroups=%d frames_per_pixel=%lld frames_per_cta=%lld smem=%zu",
                 a.nwin, pl.cl_unroll, pl.grid, kReduceBlock, pl.cluster, (long long)pl.ntiles, pl.tile4, pl.jsub, (long long)pl.ktot, (long long)pl.q, pl.cl_smem);
        return;
    }
    const ReduceVariant v = reduce_variant(a, pl);
    snprintf(buf, cap, "reduce_kernel<VEC=%d,COMP=%d,NW=%d,U=%d,MINB=%d,KG=%d> grid=%d block=%d tiles=%lld frames_per_pixel=%lld units_per_cta=%lld",
             v.vec, v.comp, v.nw, v.unroll, v.min_blocks, v.kg, pl.grid, kReduceBlock, (long long)pl.ntiles, (long long)pl.ktot, (long long)pl.q);
}

// mirrors the dispatch below (reduce_dispatch_tune / reduce_dispatch_nw)
ReduceVariant reduce_variant(const ReduceArgs &a, const ReducePlan &pl) {
    ReduceVariant v{};
    v.vec = pl.vec;
    v.comp = a.mode != SUM_SEQUENTIAL;
    v.nw = a.nwin;
    v.kg = v.comp ? pl.kg : 1;
    if (v.kg > 1) {
        v.unroll = env_int("AFK_REDUCE_KG_UNROLL", 8) == 16 ? 16 : 8;
        v.min_blocks = kReduceKgCtasPerSm;
    } else {
        const int u = reduce_tune_unroll(), c = reduce_tune_ctas();
        const bool known = (u == 8 && c == 4) || (u == 8 && c == 2) || (u == 16 && c == 2) || (u == 16 && c == 3);
        v.unroll = known ? u : 8;
        v.min_blocks = known ? c : 3;
    }
    return v;
}

cudaError_t reduce_launch(const ReduceArgs &a, const ReducePlan &pl, const ReduceWorkspace &ws, cudaStream_t s) {
    if (a.nwin < 1 || a.nwin > kMaxWindows || pl.grid < 1) return cudaErrorInvalidValue;
    if (ws.partial_bytes < pl.partial_bytes || ws.n_counters < pl.n_counters) return cudaErrorInvalidValue;
    ReduceKArgs k{};
    k.cube = a.cube;
    k.n2 = a.n2;
    k.nwin = a.nwin;
    k.woff[0] = 0;
    for (int w = 0; w < a.nwin; ++w) {
        k.start[w] = a.start[w];
        k.woff[w + 1] = k.woff[w] + (a.end[w] - a.start[w]);
        k.cnt[w] = count_as_f32(a.end[w] - a.start[w]);
    }
    for (int w = a.nwin; w < kMaxWindows; ++w) k.woff[w + 1] = k.woff[a.nwin];
    k.ktot = pl.ktot;
    k.units = pl.units;
    k.q = pl.q;
    k.ntiles = pl.ntiles;
    k.epilogue = a.epilogue;
    k.pa = a.p[0];
    k.pb = a.p[1];
    k.pfl = a.p[2];
    k.pmax = a.p[3];
    k.psign = a.ew_sign;
    if (a.n_extra_out < 0 || a.n_extra_out > kMaxPeers - 1 || a.n_signal < 0 || a.n_signal > kMaxPeers) return cudaErrorInvalidValue;
    k.out[0] = a.out;
    if (a.multicast_out && a.n_extra_out != 0) return cudaErrorInvalidValue;
    k.mc = a.multicast_out;
    for (int d = 0; d < a.n_extra_out; ++d) k.out[1 + d] = a.extra_out[d];
    k.nout = 1 + a.n_extra_out;
    k.done = a.n_signal > 0 ? a.done : nullptr;
    for (int d = 0; d < a.n_signal; ++d) k.sig[d] = a.signal[d];
    k.nsig = a.n_signal;
    k.epoch = a.epoch;
    if (a.n_range_peer < 0 || a.n_range_peer > kMaxPeers - 1) return cudaErrorInvalidValue;
    k.n_range_peer = a.n_range_peer;
    for (int d = 0; d < a.n_range_peer; ++d) k.range_peer[d] = a.range_peer[d];
    k.minmax = a.minmax;
    k.partials = static_cast<double2 *>(ws.partials);
    k.counters = ws.counters;
    k.range_out = a.range_out;
    k.range_slots = reinterpret_cast<unsigned *>(static_cast<char *>(ws.partials) + pl.range_slot_offset);
    k.range_counter = ws.counters + pl.range_counter_index;
    const bool comp = a.mode != SUM_SEQUENTIAL;
    if (!comp && pl.kg != 1) return cudaErrorInvalidValue;
    if (pl.cluster > 0) {
        if (!comp || pl.vec != 4 || a.n_signal > 0 || a.n_extra_out > 0 || pl.cl_smem > kClusterMaxSmem) return cudaErrorInvalidValue;
        k.tile4 = pl.tile4;
        k.jsub = pl.jsub;
        return reduce_cluster_dispatch(k, pl, s);
    }
    if (pl.vec == 4) return comp ? reduce_dispatch_tune<4, true>(k, pl.grid, pl.kg, s) : reduce_dispatch_tune<4, false>(k, pl.grid, 1, s);
    return comp ? reduce_dispatch_tune<1, true>(k, pl.grid, pl.kg, s) : reduce_dispatch_tune<1, false>(k, pl.grid, 1, s);
}

// ----------------------------------------------------------------------------------------------------
// K3: fused elementwise operators -- one HBM pass, float4 lanes, grid-stride
// ----------------------------------------------------------------------------------------------------
struct MapK {
    float a, b, c, ln_a;
    int flag0, flag1;
    float d, e;
    const unsigned *mm; // see MapArgs::bc_minmax
};

// powf(x, e) with a launch-uniform exponent.  The exponents a user actually sets (1 is the default of both slots, -1
// spells a ratio, 2 a square, 0 switches an input off) have exact closed forms that C99 Annex F fixes for every x, NaN,
// zeros and infinities included: pow(x, 1) = x, pow(x, 0) = 1, pow(x, -1) = 1/x and pow(x, 2) = x*x correctly rounded
// (libm's powf is within 1 ulp of them; tier B).  The branch is uniform across the grid; anything else calls powf.
__device__ __forceinline__ float pow_uniform(float x, float e) {
    if (e == 1.0f) return x;
    if (e == -1.0f) return __fdiv_rn(1.0f, x);
    if (e == 2.0f) return __fmul_rn(x, x);
    if (e == 0.0f) return 1.0f;
    return powf(x, e);
}

template <int OP> __device__ __forceinline__ float map_one(float u, float v, const MapK &k) {
    if (OP == MAP_LINCOMP) return __fadd_rn(__fmul_rn(u, k.a), __fmul_rn(v, k.b));
    if (OP == MAP_RATIO) return __fdiv_rn(u, v);
    if (OP == MAP_MULTIPLIER) return __fmul_rn(pow_uniform(u, k.a), pow_uniform(v, k.b));
    if (OP == MAP_EW) return ew_value(u, v, k.a, k.b, k.c);
    if (OP == MAP_EMISSION_LINE) {
        float t = __fsub_rn(u, v);
        t = __fmul_rn(t, k.a);
        t = __fmul_rn(t, k.c);
        return (t > k.b) ? 0.0f : t;
    }
    if (OP == MAP_LOGSCALE) {
        float d = __fdiv_rn(__fsub_rn(u, k.b), __fsub_rn(k.c, k.b));
        float t = __fadd_rn(__fmul_rn(k.a, d), 1.0f);
        return __fdiv_rn(logf(t), k.ln_a);
    }
    if (OP == MAP_VELOCITY) return __fdiv_rn(__fmul_rn(__fsub_rn(u, k.a), 3e5f), k.a); // (image - w_0) * c / w_0
    if (OP == MAP_HISTTRANS) {
        // min = a, max = b, parameters = (c, d, e): every step rounded like the closure at lib.rs:574-586
        const float range = __fsub_rn(k.b, k.a);
        const float point = __fdiv_rn(__fsub_rn(u, k.a), range);
        const float x = __fdiv_rn(__fsub_rn(point, k.c), __fsub_rn(k.e, k.c));
        float c;
        if (x > 1.0f) c = 1.0f;
        else if (x < 0.0f) c = 0.0f;
        else c = __fdiv_rn(__fmul_rn(__fsub_rn(k.d, 1.0f), x), __fsub_rn(__fmul_rn(__fsub_rn(__fmul_rn(2.0f, k.d), 1.0f), x), k.d));
        return __fadd_rn(__fmul_rn(c, range), k.a);
    }
    if (OP == MAP_LOG10) return log10f(u);
    if (OP == MAP_NEGATION) return -u;
    if (OP == MAP_NORMALIZE) return __fdiv_rn(__fsub_rn(u, k.a), __fsub_rn(k.b, k.a));
    if (OP == MAP_CLIP) return ((k.flag0 && u >= k.a) || (k.flag1 && u <= k.b)) ? __int_as_float(0x7fc00000) : u;
    if (OP == MAP_REPLACE_NAN) return (u != u) ? k.a : u;
    return u;
}

// A thread issues U independent 16-byte loads per operand before it computes (U x 2048 threads x 16 B per SM in
// flight: one load per thread covers only ~2/3 of the HBM latency-bandwidth product on B200).
template <int OP, bool BINARY>
__global__ void __launch_bounds__(256) map_kernel(const float *__restrict__ u, const float *__restrict__ v,
                                                   float *__restrict__ out, int64_t n, int vec_ok, MapK k) {
    constexpr int U = BINARY ? 2 : 4;
    // programmatic dependent launch (map_go): this grid may be scheduled while the kernel before it in the stream drains;
    // nothing of that kernel's output is touched before this point
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (OP == MAP_LOGSCALE && k.mm != nullptr) {
        // image_min_max (lib.rs:180-197) left on the device by minmax_kernel; FLT_MAX / -FLT_MAX are the operator's initial values
        const bool any = k.mm[2] != 0u;
        k.b = any ? fminf(3.402823466e38f, key_f32(~k.mm[0])) : 3.402823466e38f;
        k.c = any ? fmaxf(-3.402823466e38f, key_f32(k.mm[1])) : -3.402823466e38f;
    }
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t nthreads = (int64_t)gridDim.x * blockDim.x;
    int64_t done = 0;
    if (vec_ok) {
        const int64_t n4 = n >> 2;
        const float4 *u4 = reinterpret_cast<const float4 *>(u);
        const float4 *v4 = reinterpret_cast<const float4 *>(v);
        float4 *o4 = reinterpret_cast<float4 *>(out);
        for (int64_t i = tid; i < n4; i += nthreads * U) {
            float4 x[U], y[U];
#pragma unroll
            for (int j = 0; j < U; ++j) {
                const int64_t idx = i + j * nthreads;
                x[j] = make_float4(0, 0, 0, 0);
                y[j] = make_float4(0, 0, 0, 0);
                if (idx < n4) {
                    x[j] = ld_stream(u4 + idx);
                    if (BINARY) y[j] = ld_stream(v4 + idx);
                }
            }
#pragma unroll
            for (int j = 0; j < U; ++j) {
                const int64_t idx = i + j * nthreads;
                if (idx < n4) {
                    float4 r;
                    r.x = map_one<OP>(x[j].x, y[j].x, k);
                    r.y = map_one<OP>(x[j].y, y[j].y, k);
                    r.z = map_one<OP>(x[j].z, y[j].z, k);
                    r.w = map_one<OP>(x[j].w, y[j].w, k);
                    o4[idx] = r;
                }
            }
        }
        done = n4 << 2;
    }
    for (int64_t i = done + tid; i < n; i += nthreads) out[i] = map_one<OP>(u[i], BINARY ? v[i] : 0.0f, k);
}

static int grid_for(int64_t work_items, int block, int sm_count, int ctas_per_sm) {
    int64_t need = (work_items + block - 1) / block;
    int64_t cap = (int64_t)sm_count * ctas_per_sm;
    if (need < 1) need = 1;
    return (int)(need < cap ? need : cap);
}

// AFK_PDL=0 switches programmatic dependent launch off (plain stream order)
static bool pdl_enabled() { static const int v = env_int("AFK_PDL", 1); return v != 0; }

template <int OP, bool BINARY> static void map_go(const MapArgs &a, const MapK &k, int vec_ok, int grid, cudaStream_t s) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(256);
    cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    cudaLaunchKernelEx(&cfg, map_kernel<OP, BINARY>, a.u, a.v, a.out, a.n, vec_ok, k);
}

cudaError_t map_launch(const MapArgs &a, int sm_count, cudaStream_t s) {
    if (a.n <= 0) return cudaSuccess;
    MapK k{a.a, a.b, a.c, 0.0f, a.flag0, a.flag1, a.d, a.e, a.bc_minmax};
    if (a.op == MAP_LOGSCALE) k.ln_a = logf(a.a); // a.ln(): host libm, same call the reference makes
    auto al = [](const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; };
    const int vec_ok = al(a.u) && al(a.out) && (a.v == nullptr || al(a.v));
    const int grid = grid_for(vec_ok ? (a.n + 3) / 4 : a.n, 256, sm_count, 8);
    switch (a.op) {
    case MAP_LINCOMP: map_go<MAP_LINCOMP, true>(a, k, vec_ok, grid, s); break;
    case MAP_RATIO: map_go<MAP_RATIO, true>(a, k, vec_ok, grid, s); break;
    case MAP_MULTIPLIER: map_go<MAP_MULTIPLIER, true>(a, k, vec_ok, grid, s); break;
    case MAP_EW: map_go<MAP_EW, true>(a, k, vec_ok, grid, s); break;
    case MAP_EMISSION_LINE: map_go<MAP_EMISSION_LINE, true>(a, k, vec_ok, grid, s); break;
    case MAP_LOGSCALE: map_go<MAP_LOGSCALE, false>(a, k, vec_ok, grid, s); break;
    case MAP_LOG10: map_go<MAP_LOG10, false>(a, k, vec_ok, grid, s); break;
    case MAP_NEGATION: map_go<MAP_NEGATION, false>(a, k, vec_ok, grid, s); break;
    case MAP_NORMALIZE: map_go<MAP_NORMALIZE, false>(a, k, vec_ok, grid, s); break;
    case MAP_CLIP: map_go<MAP_CLIP, false>(a, k, vec_ok, grid, s); break;
    case MAP_REPLACE_NAN: map_go<MAP_REPLACE_NAN, false>(a, k, vec_ok, grid, s); break;
    case MAP_VELOCITY: map_go<MAP_VELOCITY, false>(a, k, vec_ok, grid, s); break;
    case MAP_HISTTRANS: map_go<MAP_HISTTRANS, false>(a, k, vec_ok, grid, s); break;
    default: return cudaErrorInvalidValue;
    }
    AFK_COUNT_LAUNCH();
    return cudaGetLastError();
}

// per-pixel operators across the 3 channels of a [3][n] image: 12 bytes read + 12 written per pixel, one pass
template <int OP> __device__ __forceinline__ void color3_one(float r, float g, float b, float beta, double asinh_beta, float (&o)[3]) {
    if (OP == C3_ARCSINH) {
        // l = (data[0] + data[1] + data[2]) / 3.0; s_fact = asinh((l * beta) as f64) / (l as f64 * asinh(beta as f64)) as f32
        const float l = __fdiv_rn(__fadd_rn(__fadd_rn(r, g), b), 3.0f);
        const float s = (float)(asinh((double)__fmul_rn(l, beta)) / ((double)l * asinh_beta));
        o[0] = __fmul_rn(s, r); o[1] = __fmul_rn(s, g); o[2] = __fmul_rn(s, b);
    } else {
        // Rust f32::max / min return the other operand when one is NaN == fmaxf / fminf
        const float mx = fmaxf(fmaxf(r, g), b), mn = fminf(fminf(r, g), b);
        const float range = __fsub_rn(mx, mn);
        float hue = 0.0f;
        if (r >= g && r >= b) hue = __fmul_rn(60.0f, __fdiv_rn(__fsub_rn(g, b), range));
        else if (g >= r && g >= b) hue = __fadd_rn(__fmul_rn(60.0f, __fdiv_rn(__fsub_rn(b, r), range)), 120.0f);
        else if (b >= r && b >= g) hue = __fadd_rn(__fmul_rn(60.0f, __fdiv_rn(__fsub_rn(r, g), range)), 240.0f);
        if (hue < 0.0f) hue = __fadd_rn(hue, 360.0f);
        o[0] = __fmul_rn(__fdiv_rn(hue, 360.0f), 65535.0f);
        o[1] = __fmul_rn(__fdiv_rn(range, mx), 65535.0f);
        o[2] = mx;
    }
}

template <int OP>
__global__ void __launch_bounds__(256) color3_kernel(const float *__restrict__ img, int64_t n, float beta, double asinh_beta,
                                                      float *__restrict__ out, int vec_ok) {
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t nthreads = (int64_t)gridDim.x * blockDim.x;
    int64_t done = 0;
    if (vec_ok) {
        const int64_t n4 = n >> 2;
        for (int64_t i = tid; i < n4; i += nthreads) {
            const float4 r = ld_stream(reinterpret_cast<const float4 *>(img) + i);
            const float4 g = ld_stream(reinterpret_cast<const float4 *>(img + n) + i);
            const float4 b = ld_stream(reinterpret_cast<const float4 *>(img + 2 * n) + i);
            float o0[3], o1[3], o2[3], o3[3];
            color3_one<OP>(r.x, g.x, b.x, beta, asinh_beta, o0);
            color3_one<OP>(r.y, g.y, b.y, beta, asinh_beta, o1);
            color3_one<OP>(r.z, g.z, b.z, beta, asinh_beta, o2);
            color3_one<OP>(r.w, g.w, b.w, beta, asinh_beta, o3);
            reinterpret_cast<float4 *>(out)[i] = make_float4(o0[0], o1[0], o2[0], o3[0]);
            reinterpret_cast<float4 *>(out + n)[i] = make_float4(o0[1], o1[1], o2[1], o3[1]);
            reinterpret_cast<float4 *>(out + 2 * n)[i] = make_float4(o0[2], o1[2], o2[2], o3[2]);
        }
        done = n4 << 2;
    }
    for (int64_t i = done + tid; i < n; i += nthreads) {
        float o[3];
        color3_one<OP>(img[i], img[n + i], img[2 * n + i], beta, asinh_beta, o);
        out[i] = o[0]; out[n + i] = o[1]; out[2 * n + i] = o[2];
    }
}

cudaError_t color3_launch(int op, const float *img, int64_t n, float beta, float *out, int sm_count, cudaStream_t s) {
    if (n <= 0) return cudaSuccess;
    const int vec_ok = (n % 4 == 0) && ((reinterpret_cast<uintptr_t>(img) & 15u) == 0) && ((reinterpret_cast<uintptr_t>(out) & 15u) == 0);
    const int grid = grid_for(vec_ok ? (n + 3) / 4 : n, 256, sm_count, 8);
    const double asinh_beta = asinh((double)beta); // host libm: the same f64 value for every pixel, as in the reference
    if (op == C3_ARCSINH) color3_kernel<C3_ARCSINH><<<grid, 256, 0, s>>>(img, n, beta, asinh_beta, out, vec_ok);
    else if (op == C3_TO_HSV) color3_kernel<C3_TO_HSV><<<grid, 256, 0, s>>>(img, n, beta, asinh_beta, out, vec_ok);
    else return cudaErrorInvalidValue;
    AFK_COUNT_LAUNCH();
    return cudaGetLastError();
}

// apply_tone_curve, lib.rs:1858-1871: out = 65535 * table[(v * ts / 65535) as usize] / ts.  Rust `as usize`: truncate,
// saturate, NaN -> 0.  The value depends on the pixel only through the table index, so the second half (one multiplication
// and one IEEE division per pixel in the reference) is evaluated once per table entry when a CTA loads the table --
// the same two roundings on the same inputs, hence the same bits.
__device__ __forceinline__ float div_uniform(float a, float s, float rs);
template <bool FAST = false>
__device__ __forceinline__ unsigned tone_key(float v, int table_len, float ts, unsigned &oob) {
    // FAST: the division by the constant 65535 as div_uniform (exact for |numerator| <= 2^100; a quotient too small for an
    // exact residual truncates to index 0 either way); anything larger, Inf and NaN take the IEEE division
    const float num = __fmul_rn(v, ts);
    float t;
    if (FAST && fabsf(num) <= 1.2676506e30f) t = div_uniform(num, 65535.0f, (float)(1.0 / 65535.0));
    else t = __fdiv_rn(num, 65535.0f);
    unsigned key;
    if (!(t > 0.0f)) key = 0u;                    // NaN, negatives, zero
    else if (t >= 4294967296.0f) key = 0xFFFFFFFFu; // saturate
    else key = (unsigned)t;                         // cvt.rzi
    if (key >= (unsigned)table_len) { // the reference panics (index out of bounds); flag it, stay in bounds
        oob = 1u;
        key = (unsigned)table_len - 1u;
    }
    return key;
}
__device__ __forceinline__ float tone_value(float entry, float ts) { return __fdiv_rn(__fmul_rn(65535.0f, entry), ts); }

template <bool FAST>
__global__ void __launch_bounds__(256) tone_kernel(const float *__restrict__ img, int64_t n, const float *__restrict__ table,
                                                    int table_len, float *__restrict__ out, unsigned *oob_flag, int vec_ok) {
    extern __shared__ float s_table[]; // the OUTPUT value of every table index
    const float ts = (float)(table_len - 1);
    for (int i = threadIdx.x; i < table_len; i += blockDim.x) s_table[i] = tone_value(table[i], ts);
    __syncthreads();
    unsigned oob = 0u;
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t nthreads = (int64_t)gridDim.x * blockDim.x;
    int64_t done = 0;
    if (vec_ok) {
        const int64_t n4 = n >> 2;
        const float4 *i4 = reinterpret_cast<const float4 *>(img);
        float4 *o4 = reinterpret_cast<float4 *>(out);
        for (int64_t i = tid; i < n4; i += nthreads) {
            float4 x = ld_stream(i4 + i), r;
            r.x = s_table[tone_key<FAST>(x.x, table_len, ts, oob)];
            r.y = s_table[tone_key<FAST>(x.y, table_len, ts, oob)];
            r.z = s_table[tone_key<FAST>(x.z, table_len, ts, oob)];
            r.w = s_table[tone_key<FAST>(x.w, table_len, ts, oob)];
            o4[i] = r;
        }
        done = n4 << 2;
    }
    for (int64_t i = done + tid; i < n; i += nthreads) out[i] = s_table[tone_key<FAST>(img[i], table_len, ts, oob)];
    if (oob) *reinterpret_cast<volatile unsigned *>(oob_flag) = 1u; // plain store: the word may be mapped host memory
}

cudaError_t tone_curve_launch(const float *img, int64_t n, const float *table_dev, int table_len, float *out,
                              unsigned *oob_flag, int sm_count, cudaStream_t s) {
    if (table_len < 1 || table_len > kMaxToneTable) return cudaErrorInvalidValue;
    if (n <= 0) return cudaSuccess;
    auto al = [](const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; };
    const int vec_ok = al(img) && al(out);
    const int grid = grid_for(vec_ok ? (n + 3) / 4 : n, 256, sm_count, 8);
    if (display_exact()) tone_kernel<false><<<grid, 256, (size_t)table_len * sizeof(float), s>>>(img, n, table_dev, table_len, out, oob_flag, vec_ok);
    else tone_kernel<true><<<grid, 256, (size_t)table_len * sizeof(float), s>>>(img, n, table_dev, table_len, out, oob_flag, vec_ok);
    AFK_COUNT_LAUNCH();
    return cudaGetLastError();
}

// ----------------------------------------------------------------------------------------------------
// K4: min/max ignoring NaN (image_min_max lib.rs:180-197; lims::get_vmin/get_vmax lims.rs:7-35)
// ----------------------------------------------------------------------------------------------------
// host_out (may be null): mapped pinned host words.  The last CTA to flush (ticket = result[3]) stores the three final words
// there with plain stores -- the host needs a stream synchronisation and no copy -- and returns result[0..3] to zero, so the
// same four device words serve the next launch on the stream without a memset.
__global__ void __launch_bounds__(256) minmax_kernel(const float *__restrict__ img, int64_t n, unsigned *result, int vec_ok, unsigned *host_out) {
    // running values start at +-Inf so that +-Inf inputs are seen; the FLT_MAX init of image_min_max is applied
    // by the host when it decodes the result
    float mn = __int_as_float(0x7f800000), mx = __int_as_float(0xff800000);
    bool any = false;
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t nthreads = (int64_t)gridDim.x * blockDim.x;
    int64_t done = 0;
    if (vec_ok) {
        // the compiler unrolls this loop four times with the loads ahead; an explicit batch of 4 loads measured slower (21 vs 15 us)
        const int64_t n4 = n >> 2;
        const float4 *i4 = reinterpret_cast<const float4 *>(img);
        for (int64_t i = tid; i < n4; i += nthreads) {
            float4 x = ld_stream(i4 + i);
            float e[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
            for (int j = 0; j < 4; ++j)
                if (e[j] == e[j]) { mn = fminf(mn, e[j]); mx = fmaxf(mx, e[j]); any = true; }
        }
        done = n4 << 2;
    }
    for (int64_t i = done + tid; i < n; i += nthreads) {
        float e = img[i];
        if (e == e) { mn = fminf(mn, e); mx = fmaxf(mx, e); any = true; }
    }
    block_minmax_flush(mn, mx, any, result);
    if (host_out && threadIdx.x == 0) { // thread 0 issued this CTA's three atomics
        __threadfence();
        if (atomicAdd(result + 3, 1u) == gridDim.x - 1u) {
            __threadfence();
            const unsigned a = __ldcg(result + 0), b = __ldcg(result + 1), c = __ldcg(result + 2);
            volatile unsigned *h = host_out;
            h[0] = a;
            h[1] = b;
            h[2] = c;
            result[0] = 0u;
            result[1] = 0u;
            result[2] = 0u;
            result[3] = 0u;
        }
    }
}

static cudaError_t minmax_go(const float *img, int64_t n, unsigned *result, unsigned *host_out, int sm_count, cudaStream_t s) {
    if (n <= 0) return cudaSuccess;
    const int vec_ok = (reinterpret_cast<uintptr_t>(img) & 15u) == 0;
    const int grid = grid_for(vec_ok ? (n + 3) / 4 : n, 256 * 4, sm_count, 8);
    minmax_kernel<<<grid, 256, 0, s>>>(img, n, result, vec_ok, host_out);
    AFK_COUNT_LAUNCH();
    return cudaGetLastError();
}
cudaError_t minmax_launch(const float *img, int64_t n, unsigned *result, int sm_count, cudaStream_t s) {
    return minmax_go(img, n, result, nullptr, sm_count, s);
}
// words4: four device words that are zero when the launch starts and zero again when it ends (self-resetting; word 3 is the
// ticket); host_out: three mapped pinned host words that hold {~key(min), key(max), any} once the stream is synchronised
cudaError_t minmax_launch_readback(const float *img, int64_t n, unsigned *words4, unsigned *host_out, int sm_count, cudaStream_t s) {
    if (!words4 || !host_out) return cudaErrorInvalidValue;
    return minmax_go(img, n, words4, host_out, sm_count, s);
}

// ----------------------------------------------------------------------------------------------------
// K6: down_sampling = decimation (lib.rs:1894-1962): out[z][j][i] = img[z][j*n][i*n]
// ----------------------------------------------------------------------------------------------------
// IDX = unsigned when every index fits in 32 bits (the usual case): 64-bit div/mod is ~4x the instructions
template <typename IDX>
__global__ void __launch_bounds__(256) downsample_kernel(const float *__restrict__ img, int64_t ny_, int64_t nx_, int64_t n_,
                                                          float *__restrict__ out, int64_t oy_, int64_t ox_, int64_t total_) {
    const IDX ny = (IDX)ny_, nx = (IDX)nx_, n = (IDX)n_, oy = (IDX)oy_, ox = (IDX)ox_, total = (IDX)total_;
    const IDX step = (IDX)gridDim.x * blockDim.x;
    for (IDX i = (IDX)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += step) {
        const IDX t = i / ox, x = i - t * ox, z = t / oy, y = t - z * oy;
        out[i] = img[(size_t)((z * ny + y * n) * nx + x * n)];
    }
}

cudaError_t downsample_launch(const float *img, int64_t c, int64_t ny, int64_t nx, int64_t n, float *out, int64_t oy,
                              int64_t ox, cudaStream_t s) {
    const int64_t total = c * oy * ox;
    if (total <= 0) return cudaSuccess;
    int64_t blocks = (total + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    // 32-bit indices when the input, the output and the grid-stride overshoot all stay below 2^32
    const bool small = c * ny * nx < (int64_t(1) << 31) && total < (int64_t(1) << 31);
    if (small) downsample_kernel<unsigned><<<(int)blocks, 256, 0, s>>>(img, ny, nx, n, out, oy, ox, total);
    else downsample_kernel<int64_t><<<(int)blocks, 256, 0, s>>>(img, ny, nx, n, out, oy, ox, total);
    AFK_COUNT_LAUNCH();
    return cudaGetLastError();
}

// ----------------------------------------------------------------------------------------------------
// K5: colour mapping (aflak_plot/src/imshow/lut.rs:132-160, image.rs:20-84, util.rs:3-14)
// ----------------------------------------------------------------------------------------------------
struct CmapK {
    float vmin[3], vmax[3], l0[3], l1[3], l2[3];
    int chain;
    float ls_a, ls_min, ls_max, ls_lna, nm_min, nm_max;
    int tone_len;
    // plain colour map, launch-uniform divisors with host-checked ranges (see div_uniform): reciprocals of vmax - vmin and l2 - l0
    int fast_div;
    float rs_range, rs_lims;
    // fused chain, guard-band evaluation of the tone index (see chain_key_fast)
    int fast_chain;
    float fc_k2, fc_off, fc_delta;
    unsigned fc_lo_bits, fc_span_bits;
};

// a / s for a launch-uniform divisor s whose correctly rounded reciprocal rs = RN(1/s) was computed once (on the host in f64:
// the quotient of two f32 is never within 2^-49 relative of a 25-bit midpoint, so rounding the f64 quotient to f32 is the
// correctly rounded f32 reciprocal).  Markstein's theorem: with y = RN(1/b) and q0 a faithful quotient, q0 + RN(a - b q0) y
// rounds to RN(a/b) as long as nothing overflows or underflows -- three instructions instead of the ~12 of div.rn.f32 and its
// FCHK-guarded fallback.  The CALLER guarantees the range (|a| <= 2^80 and a quotient that is either >= 2^-100 in magnitude or
// irrelevant below that, e.g. because only a truncated index is taken from it); NaN propagates.
__device__ __forceinline__ float div_uniform(float a, float s, float rs) {
    const float q0 = __fmul_rn(a, rs);
    const float r = __fmaf_rn(-q0, s, a);
    return __fmaf_rn(r, rs, q0);
}

// mtf, lut.rs:152-160
__device__ __forceinline__ float mtf(float m, float x) {
    if (x >= 1.0f) return 1.0f;
    if (x <= 0.0f) return 0.0f;
    float num = __fmul_rn(__fsub_rn(m, 1.0f), x);
    float den = __fsub_rn(__fmul_rn(__fsub_rn(__fmul_rn(2.0f, m), 1.0f), x), m);
    return __fdiv_rn(num, den);
}

// color_at_bounds + color_at (index part), lut.rs:132-149.
// Two launch-uniform special cases drop divisions without changing the index:
//   l1 == 0.5 (the default midtone): mtf(0.5, c) = (-0.5 c) / (0 c - 0.5) = c for every c in (0, 1) -- scaling by a power of two is
//     exact; a denormal c can lose its last bit on the way, but then c * 65535 < 1 and the index is 0 either way;
//   l0 == 0 and l2 == 1 (the default limits): (x - 0) / (1 - 0) = x exactly.
// With the reference's default limits (0, 0.5, 1) one IEEE division per pixel is left instead of three.
__device__ __forceinline__ unsigned lut_index(float val, float vmin, float vmax, float l0, float l1, float l2) {
    float p = val;
    if (p < vmin) p = vmin;       // util::clamp: NaN fails both comparisons and passes through
    else if (p > vmax) p = vmax;
    float x = __fdiv_rn(__fsub_rn(p, vmin), __fsub_rn(vmax, vmin));
    float c = (l0 == 0.0f && l2 == 1.0f) ? x : __fdiv_rn(__fsub_rn(x, l0), __fsub_rn(l2, l0));
    float m;
    if (l1 == 0.5f) m = c >= 1.0f ? 1.0f : (c <= 0.0f ? 0.0f : c);
    else m = mtf(l1, c);
    float i = __fmul_rn(m, 65535.0f);
    if (i < 0.0f) i = 0.0f;
    unsigned idx;
    if (!(i > 0.0f)) idx = 0u; // NaN as usize == 0
    else if (i >= 65535.0f) idx = 65535u;
    else idx = (unsigned)i;
    return idx;
}

// lut_index with the two launch-uniform divisions done by div_uniform.  Preconditions checked by the host (colormap_launch):
// vmin < vmax, both finite with a range in [2^-60, 2^60]; 0 <= l0 < l2 <= 1 with l2 - l0 >= 2^-20.  Then p - vmin lies in
// [0, vmax - vmin] and x in [0, 1]: no overflow, and an x or c so small that the residual underflows gives index 0 either way.
// A NaN pixel takes index 0 in the reference (clamp passes it through, `NaN as usize` is 0); so does vmin (x = 0, c <= 0, m = 0):
// the clamp may therefore be the two-instruction fmax/fmin form that maps NaN to vmin.
__device__ __forceinline__ unsigned lut_index_fast(float val, float vmin, float vmax, float l0, float l1, float l2, float rs_range, float rs_lims) {
    const float p = fminf(fmaxf(val, vmin), vmax);
    const float x = div_uniform(__fsub_rn(p, vmin), __fsub_rn(vmax, vmin), rs_range);
    const float c = (l0 == 0.0f && l2 == 1.0f) ? x : div_uniform(__fsub_rn(x, l0), __fsub_rn(l2, l0), rs_lims);
    float m;
    if (l1 == 0.5f) m = fminf(fmaxf(c, 0.0f), 1.0f);
    else m = mtf(l1, c);
    const float i = __fmul_rn(m, 65535.0f);
    unsigned idx = (i > 0.0f) ? (unsigned)i : 0u;
    return min(idx, 65535u);
}

// fused chain, per pixel: convert_to_logscale (lib.rs:1598-1611) -> normalize_image (lib.rs:269-273) -> the index half of
// apply_tone_curve (lib.rs:1858-1871).  Everything downstream of that index -- the tone curve's output value, clamp, the two
// range divisions, mtf and the colour-table look-up -- is a function of the index alone and is evaluated once per tone-table
// entry when the CTA starts (s_color), with the reference's own sequence of roundings.
__device__ __forceinline__ unsigned chain_key(float v, const CmapK &k, unsigned &oob) {
    float d = __fdiv_rn(__fsub_rn(v, k.ls_min), __fsub_rn(k.ls_max, k.ls_min));
    float t = __fdiv_rn(logf(__fadd_rn(__fmul_rn(k.ls_a, d), 1.0f)), k.ls_lna);
    t = __fdiv_rn(__fsub_rn(t, k.nm_min), __fsub_rn(k.nm_max, k.nm_min));
    return tone_key(t, k.tone_len, (float)(k.tone_len - 1), oob);
}

// Guard-band evaluation of chain_key.  Only the INDEX must equal the reference's, not the floats on the way: the index is
// trunc(kf) of a quantity kf that is, in exact arithmetic, F(t) = ((ln t / ln a - nm_min) / (nm_max - nm_min)) ts / 65535 with
// t = a d + 1.  The fast path evaluates F with one subtraction, one multiplication by the reciprocal of the range, one FFMA, one
// MUFU.LG2 and one FFMA (kf' = log2(t') k2 + off), and the host bounds |kf' - kf_reference| by delta (error analysis in
// chain_fast_setup: every rounding of the reference's nine steps, logf's 1 ulp, lg2.approx's 2^-22 + 1 ulp, the roundings of
// the constants, doubled).  When kf' is farther than delta from every integer, trunc(kf') IS the reference's index; otherwise
// -- about 2 delta of the pixels, ~0.1 % -- and for every pixel outside the domain the bound was derived for (kf' outside
// [lo, hi): t < 1, NaN, Inf, indices at or beyond the end of the table) the reference's own sequence is evaluated.
// Returns true when the pixel needs the exact path.  Rounding to the nearest integer with the 1.5 * 2^23 trick keeps
// everything on the FMA pipe (no F2I / I2F).
__device__ __forceinline__ bool chain_key_fast(float v, float ls_min, float rs, const CmapK &k, unsigned &key) {
    const float x = __fsub_rn(v, ls_min);
    const float t = __fmaf_rn(k.ls_a, __fmul_rn(x, rs), 1.0f);
    float lg;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(lg) : "f"(t));
    const float kf = __fmaf_rn(lg, k.fc_k2, k.fc_off);
    const float y = __fadd_rn(kf, 12582912.0f);          // 1.5 * 2^23: ulp 1, the low mantissa bits hold round(kf)
    const float df = __fsub_rn(kf, __fsub_rn(y, 12582912.0f)); // signed distance to the nearest integer, exact
    key = (__float_as_uint(y) & 0x3FFFFFu) - (df < 0.0f ? 1u : 0u);
    const bool in_domain = (__float_as_uint(kf) - k.fc_lo_bits) < k.fc_span_bits; // lo <= kf < hi for non-negative floats; NaN fails
    return !in_domain || fabsf(df) < k.fc_delta;
}

// One group = 4 consecutive pixels of an (input) row: map them to packed colours.
template <bool RGB3, bool CHAIN, bool VEC>
__device__ __forceinline__ void colormap_group(const float *__restrict__ img, size_t n2, size_t src_off, int cnt, const uint32_t *__restrict__ lut,
                                               const uint32_t *s_color, const CmapK &k, bool fast_chain, float fc_rs, unsigned &oob,
                                               uint32_t (&px)[4]) {
#pragma unroll
    for (int c = 0; c < (RGB3 ? 3 : 1); ++c) {
        const float *src = img + (size_t)c * n2 + src_off;
        float v[4];
        if (VEC) {
            float4 t = ld_stream(reinterpret_cast<const float4 *>(src));
            v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
        } else {
#pragma unroll
            for (int j = 0; j < 4; ++j) v[j] = j < cnt ? src[j] : 0.0f;
        }
        unsigned ckey[4] = {0u, 0u, 0u, 0u};
        if (CHAIN) {
            if (fast_chain) {
                bool slow[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) slow[j] = chain_key_fast(v[j], k.ls_min, fc_rs, k, ckey[j]);
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (slow[j]) ckey[j] = (v[j] == v[j]) ? chain_key(v[j], k, oob) : 0u; // a NaN pixel takes index 0 (NaN as usize)
            } else {
#pragma unroll
                for (int j = 0; j < 4; ++j) ckey[j] = chain_key(v[j], k, oob);
            }
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            uint32_t e;
            if (CHAIN) {
                e = s_color[ckey[j]];
            } else {
                const unsigned idx = (!RGB3 && k.fast_div) ? lut_index_fast(v[j], k.vmin[0], k.vmax[0], k.l0[0], k.l1[0], k.l2[0], k.rs_range, k.rs_lims)
                                                           : lut_index(v[j], k.vmin[c], k.vmax[c], k.l0[c], k.l1[c], k.l2[c]);
                e = __ldg(lut + (size_t)c * 65536 + idx);
            }
            if (RGB3) {
                // make_raw_image_rgb (image.rs:66-72): channel c takes component c of its own table
                const uint32_t comp = (e >> (8 * c)) & 0xFFu;
                px[j] = (c == 0 ? 0xFF000000u : px[j]) | (comp << (8 * c));
            } else {
                px[j] = e;
            }
        }
    }
}

template <int CH, bool VEC> __device__ __forceinline__ void colormap_store(uint8_t *dst, const uint32_t (&px)[4], int cnt) {
    if (CH == 4) {
        if (VEC) {
            *reinterpret_cast<uint4 *>(dst) = make_uint4(px[0], px[1], px[2], px[3]);
        } else {
            for (int j = 0; j < cnt; ++j) {
                dst[4 * j + 0] = px[j] & 0xFF; dst[4 * j + 1] = (px[j] >> 8) & 0xFF;
                dst[4 * j + 2] = (px[j] >> 16) & 0xFF; dst[4 * j + 3] = 0xFF;
            }
        }
    } else {
        if (VEC) {
            // 4 px = 12 bytes: r0 g0 b0 r1 | g1 b1 r2 g2 | b2 r3 g3 b3
            const uint32_t a0 = px[0] & 0xFFFFFFu, a1 = px[1] & 0xFFFFFFu, a2 = px[2] & 0xFFFFFFu, a3 = px[3] & 0xFFFFFFu;
            uint32_t *d32 = reinterpret_cast<uint32_t *>(dst);
            d32[0] = a0 | (a1 << 24);
            d32[1] = (a1 >> 8) | (a2 << 16);
            d32[2] = (a2 >> 16) | (a3 << 8);
        } else {
            for (int j = 0; j < cnt; ++j) {
                dst[3 * j + 0] = px[j] & 0xFF; dst[3 * j + 1] = (px[j] >> 8) & 0xFF; dst[3 * j + 2] = (px[j] >> 16) & 0xFF;
            }
        }
    }
}

// VEC: nx % 4 == 0, 16-byte aligned image and output, fewer than 2^31 pixels -- the case that matters for throughput.  The
// grid-stride loop then carries (row, group) along by addition instead of dividing the flat index every iteration and uses
// 32-bit offsets: the loop overhead was ~100 of the ~190 instructions per 4 pixels (profiles/r2_c4_kernels_ncu_full.json).
template <bool RGB3, bool CHAIN, int CH, bool VEC>
__global__ void __launch_bounds__(256) colormap_kernel(const float *__restrict__ img, int64_t ny, int64_t nx,
                                                        const uint32_t *__restrict__ lut, const float *__restrict__ tone,
                                                        int flip_y, uint8_t *__restrict__ out, unsigned *oob_flag,
                                                        const unsigned *__restrict__ ls_minmax, CmapK k) {
    extern __shared__ uint32_t s_color[]; // CHAIN: the packed colour of every tone-table index
    if (CHAIN) {
        if (ls_minmax != nullptr) {
            // image_min_max (lib.rs:180-197) left on the device by minmax_kernel: {~key(min), key(max), any}; the operator's
            // FLT_MAX / -FLT_MAX initial values bound the result
            const bool any = ls_minmax[2] != 0u;
            k.ls_min = any ? fminf(3.402823466e38f, key_f32(~ls_minmax[0])) : 3.402823466e38f;
            k.ls_max = any ? fmaxf(-3.402823466e38f, key_f32(ls_minmax[1])) : -3.402823466e38f;
        }
        const float ts = (float)(k.tone_len - 1);
        for (int i = threadIdx.x; i < k.tone_len; i += blockDim.x)
            s_color[i] = __ldg(lut + lut_index(tone_value(tone[i], ts), k.vmin[0], k.vmax[0], k.l0[0], k.l1[0], k.l2[0]));
        __syncthreads();
    }
    // fast chain: the range is only known here when image_min_max stayed on the device; its reciprocal is one IEEE
    // reciprocal per thread, and a degenerate range (constant image, Inf) falls back to the exact sequence
    float fc_rs = 0.0f;
    bool fast_chain = false;
    if (CHAIN && k.fast_chain) {
        const float range = __fsub_rn(k.ls_max, k.ls_min);
        fast_chain = range >= 7.8886e-31f && range <= 1.2676e30f && fabsf(k.ls_min) <= 1.2676e30f; // 2^-100 .. 2^100
        fc_rs = __frcp_rn(range);
    }
    unsigned oob = 0u;
    const size_t n2 = (size_t)(ny * nx);
    if (VEC) {
        const unsigned gpr = (unsigned)(nx >> 2), rows = (unsigned)ny, total = rows * gpr;
        const unsigned stride = gridDim.x * blockDim.x;
        const unsigned sy = stride / gpr, sx = stride - sy * gpr;
        unsigned gi = blockIdx.x * blockDim.x + threadIdx.x;
        unsigned y = gi / gpr, xg = gi - y * gpr;
        for (; gi < total; gi += stride) {
            const unsigned ys = flip_y ? rows - 1u - y : y;
            uint32_t px[4];
            colormap_group<RGB3, CHAIN, true>(img, n2, (size_t)(ys * (unsigned)nx + (xg << 2)), 4, lut, s_color, k, fast_chain, fc_rs, oob, px);
            colormap_store<CH, true>(out + (size_t)gi * (4 * CH), px, 4);
            xg += sx;
            y += sy;
            if (xg >= gpr) { xg -= gpr; ++y; }
        }
    } else {
        const int64_t groups_per_row = (nx + 3) >> 2;
        const int64_t total = ny * groups_per_row;
        for (int64_t gi = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; gi < total; gi += (int64_t)gridDim.x * blockDim.x) {
            const int64_t y = gi / groups_per_row;
            const int64_t x0 = (gi - y * groups_per_row) << 2;
            const int64_t ys = flip_y ? (ny - 1 - y) : y;
            const int cnt = (int)min((int64_t)4, nx - x0);
            uint32_t px[4];
            colormap_group<RGB3, CHAIN, false>(img, n2, (size_t)(ys * nx + x0), cnt, lut, s_color, k, fast_chain, fc_rs, oob, px);
            colormap_store<CH, false>(out + (size_t)(y * nx + x0) * CH, px, cnt);
        }
    }
    if (CHAIN && oob) *reinterpret_cast<volatile unsigned *>(oob_flag) = 1u; // plain store: the word may be mapped host memory
}

static bool fast_paths_enabled() { return !display_exact(); }

static float f32_reciprocal(float s) { return (float)(1.0 / (double)s); } // correctly rounded, see div_uniform

// plain colour map: are the two launch-uniform divisors in the range div_uniform needs?
static void lut_fast_setup(const ColormapArgs &a, CmapK &k) {
    k.fast_div = 0;
    if (a.rgb3 || a.chain || !fast_paths_enabled()) return;
    const float vmin = a.vmin[0], vmax = a.vmax[0], l0 = a.lims[0][0], l2 = a.lims[0][2];
    const float range = vmax - vmin, lr = l2 - l0;
    if (!(vmin < vmax) || !(fabsf(vmin) <= 1.1529215e18f) || !(fabsf(vmax) <= 1.1529215e18f)) return; // 2^60
    if (!(range >= 8.6736174e-19f) || !(range <= 1.1529215e18f)) return;                                // 2^-60 .. 2^60
    if (!(l0 >= 0.0f) || !(l2 <= 1.0f) || !(lr >= 9.5367432e-7f)) return;                                // 2^-20
    k.fast_div = 1;
    k.rs_range = f32_reciprocal(range);
    k.rs_lims = f32_reciprocal(lr);
}

// fused chain: constants and error bound of chain_key_fast.  Everything below is f64 arithmetic on launch-uniform values.
static void chain_fast_setup(const ColormapArgs &a, CmapK &k) {
    k.fast_chain = 0;
    if (!a.chain || !fast_paths_enabled()) return;
    const double A = a.ls_a, LNA = k.ls_lna, NMIN = a.nm_min, R = (double)(a.nm_max - a.nm_min);
    const int len = a.tone_len;
    if (!(A > 1.0) || !(A <= 1e6) || !(LNA > 0.0) || !(R > 0.0) || !std::isfinite(R) || !std::isfinite(NMIN) || len < 2) return;
    const double TS = len - 1, C = TS / 65535.0;
    const double Kd = C / (LNA * R);              // d kf / d ln t
    const double K2 = Kd * 0.69314718055994530942; // per log2
    const double OFF = -NMIN / R * C;
    const double T_HI = 2.0 * (A + 1.0);          // d <= 1 gives t <= a + 1; the domain check is on kf, see hi below
    const double Lmax = log(T_HI), lgmax = log2(T_HI);
    const double eps = ldexp(1.0, -24);
    auto ulp = [](double v) { return ldexp(1.0, (int)floor(log2(fabs(v))) - 23); };
    // (1) the reference's own roundings after t: logf (1 ulp), / ln a, - nm_min, / range, * ts, / 65535
    const double Umax = Lmax / LNA, Wmax = Umax + fabs(NMIN), Nmax = Wmax / R, Kmax = Nmax * C;
    const double e_ref = ulp(Lmax) * Kd + eps * ((Umax + Wmax) / R * C + 3.0 * Kmax);
    // (2) the fast path after t': lg2.approx (2^-22 on the mantissa's log + the rounding of the sum), the rounding of k2 and off, the FFMA
    const double e_fast = (ldexp(1.0, -22) + ulp(lgmax)) * fabs(K2) + eps * (lgmax * fabs(K2) + fabs(OFF) + Kmax + fabs(OFF));
    // (3) t' against the reference's t: x is the same subtraction; d = fl(x / S) against fl(x * RN(1/S)), fl(fl(a d) + 1) against
    //     fma(a, d', 1): at most 6 half-ulps relative for t >= 1, and d ln t = dt / t
    const double e_t = Kd * 6.0 * eps;
    const double delta = 2.0 * (e_ref + e_fast + e_t);
    if (!(delta < 0.125) || !std::isfinite(delta) || !std::isfinite(K2) || !std::isfinite(OFF)) return;
    const double lo = OFF > 0.0 ? OFF : 0.0;                       // kf' >= lo implies t' >= 1 (up to the error already in delta)
    double hi = Kd * Lmax + OFF;                                   // t' <= T_HI
    if (hi > (double)len - 2.0 * delta) hi = (double)len - 2.0 * delta; // kf' < hi implies the reference's index < len: no out-of-bounds case
    if (!(hi > lo) || hi >= 4194304.0) return;
    const float lo_f = (float)lo, hi_f = nextafterf((float)hi, 0.0f);
    if (!(hi_f > lo_f)) return;
    unsigned lo_bits, hi_bits;
    memcpy(&lo_bits, &lo_f, 4);
    memcpy(&hi_bits, &hi_f, 4);
    k.fc_k2 = (float)K2;
    k.fc_off = (float)OFF;
    k.fc_delta = (float)delta;
    k.fc_lo_bits = lo_bits;
    k.fc_span_bits = hi_bits - lo_bits;
    k.fast_chain = 1;
}

cudaError_t colormap_launch(const ColormapArgs &a, int sm_count, cudaStream_t s) {
    if (a.ny <= 0 || a.nx <= 0) return cudaSuccess;
    if (a.channels != 3 && a.channels != 4) return cudaErrorInvalidValue;
    if (a.chain && (a.rgb3 || a.tone_len < 1 || a.tone_len > kMaxToneTable)) return cudaErrorInvalidValue;
    CmapK k{};
    for (int c = 0; c < 3; ++c) {
        k.vmin[c] = a.vmin[c]; k.vmax[c] = a.vmax[c];
        k.l0[c] = a.lims[c][0]; k.l1[c] = a.lims[c][1]; k.l2[c] = a.lims[c][2];
    }
    k.chain = a.chain;
    k.ls_a = a.ls_a; k.ls_min = a.ls_min; k.ls_max = a.ls_max; k.ls_lna = logf(a.ls_a);
    k.nm_min = a.nm_min; k.nm_max = a.nm_max;
    k.tone_len = a.tone_len;
    lut_fast_setup(a, k);
    chain_fast_setup(a, k);
    const int64_t groups = a.ny * ((a.nx + 3) / 4);
    const int grid = grid_for(groups, 256, sm_count, 8);
    const bool vec = (a.nx % 4 == 0) && ((reinterpret_cast<uintptr_t>(a.img) & 15u) == 0) && ((reinterpret_cast<uintptr_t>(a.out) & 15u) == 0) &&
                     a.ny * a.nx < (int64_t(1) << 31);
    const size_t smem = a.chain ? (size_t)a.tone_len * sizeof(float) : 0;
#define AFK_CMAP(R, C, CH)                                                                                                                        \
    do {                                                                                                                                          \
        if (vec) colormap_kernel<R, C, CH, true><<<grid, 256, smem, s>>>(a.img, a.ny, a.nx, a.lut, a.tone_table, a.flip_y, a.out, a.oob_flag, a.ls_minmax, k);  \
        else colormap_kernel<R, C, CH, false><<<grid, 256, smem, s>>>(a.img, a.ny, a.nx, a.lut, a.tone_table, a.flip_y, a.out, a.oob_flag, a.ls_minmax, k); \
    } while (0)
    if (a.rgb3) {
        if (a.channels == 4) AFK_CMAP(true, false, 4); else AFK_CMAP(true, false, 3);
    } else if (a.chain) {
        if (a.channels == 4) AFK_CMAP(false, true, 4); else AFK_CMAP(false, true, 3);
    } else {
        if (a.channels == 4) AFK_CMAP(false, false, 4); else AFK_CMAP(false, false, 3);
    }
#undef AFK_CMAP
    AFK_COUNT_LAUNCH();
    return cudaGetLastError();
}

// ----------------------------------------------------------------------------------------------------
// Synthetic cube generator: bit-identical to oracle/aflak_oracle.c orc_gen_voxel (only + - * / in f32)
// ----------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t mix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

__device__ __forceinline__ float gen_voxel(uint64_t seed, int64_t nz, int64_t ny, int64_t nx, int64_t k, int64_t y, int64_t x,
                                           int pathology) {
    const uint64_t idx = ((uint64_t)k * (uint64_t)ny + (uint64_t)y) * (uint64_t)nx + (uint64_t)x;
    const uint64_t h = mix64(seed + idx);
    const float u1 = __fmul_rn(__uint2float_rn((uint32_t)(h >> 40)), 5.9604644775390625e-08f);
    const float u2 = __fmul_rn(__uint2float_rn((uint32_t)((h >> 16) & 0xFFFFFFu)), 5.9604644775390625e-08f);
    const float ph = __fmul_rn(__uint2float_rn((uint32_t)((uint64_t)(x * 3 + y * 5) & 255u)), 0.00390625f);
    const float fnz = __ll2float_rn(nz);
    float k0 = __fsub_rn(ph, 0.5f);
    k0 = __fmul_rn(k0, 0.125f);
    k0 = __fadd_rn(k0, 0.5f);
    k0 = __fmul_rn(k0, fnz);
    float sigma = __fadd_rn(__fmul_rn(fnz, 0.0078125f), 2.0f);
    float d = __fdiv_rn(__fsub_rn(__ll2float_rn(k), k0), sigma);
    float dd = __fadd_rn(__fmul_rn(d, d), 1.0f);
    float line = __fdiv_rn(2.0f, dd);
    float cont = __fadd_rn(__fmul_rn(0.1f, u1), 1.0f);
    float noise = __fmul_rn(__fsub_rn(u2, 0.5f), 0.05f);
    float v = __fadd_rn(__fadd_rn(cont, line), noise);
    if (pathology) {
        const uint32_t r = (uint32_t)(h & 0xFFFFu);
        if (r < 16u) v = __int_as_float(0x7fc00000);
        else if (r < 24u) v = __int_as_float(0x7f800000);
        else if (r < 32u) v = __int_as_float(0xff800000);
        else if (r < 48u) v = 0.0f;
        else if (r < 64u) v = -v;
        else if (r < 80u) v = 1e-40f;
    }
    return v;
}

__global__ void __launch_bounds__(256) gen_cube_kernel(uint64_t seed, int64_t nz, int64_t ny, int64_t nx, int64_t y0, int64_t rows,
                                                        int pathology, float *__restrict__ out, int64_t total) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t x = i % nx, t = i / nx, r = t % rows, k = t / rows;
        out[i] = gen_voxel(seed, nz, ny, nx, k, y0 + r, x, pathology);
    }
}

cudaError_t gen_cube_launch(uint64_t seed, int64_t nz, int64_t ny, int64_t nx, int64_t y0, int64_t rows, int pathology,
                            float *out, cudaStream_t s) {
    const int64_t total = nz * rows * nx;
    if (total <= 0) return cudaSuccess;
    int64_t blocks = (total + 255) / 256;
    if (blocks > 148 * 32) blocks = 148 * 32;
    gen_cube_kernel<<<(int)blocks, 256, 0, s>>>(seed, nz, ny, nx, y0, rows, pathology, out, total);
    AFK_COUNT_LAUNCH();
    return cudaGetLastError();
}

} // namespace afk
