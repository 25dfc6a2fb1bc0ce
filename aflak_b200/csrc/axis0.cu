// axis0.cu -- the other per-spaxel reductions along the spectral axis: variance / stddev, the
// max/min fold, argmax/argmin wavelength maps, centrobaric wavelength (with and without per-pixel masks) and the
// per-pixel median.  Same access pattern as the integral (kernels.cu): a thread owns 4 consecutive spaxels and
// streams frames with 8 independent 16-byte loads in flight; every voxel is read from HBM exactly once.
//
// Decomposition: grid = (pixel tiles of 256*VEC spaxels) x (ksplit chunks of the spectral range).  ksplit == 1
// (any image with >= 2 waves of tiles) finishes in registers.  Small images split the spectral range so the chip
// is filled; the per-chunk states go through scratch and a second tiny kernel merges them in chunk order
// (deterministic).  Operators whose reference arithmetic is order dependent and not mergeable bit-exactly
// (Welford, sequential sums) are never split in SUM_SEQUENTIAL mode.
//
// Reference citations are relative to /root/reference/src/.
#include <cstdlib>
#include "kernels.h"

#include <cstdio>

#include <cfloat>
#include <cstdint>

namespace afk {

__device__ __forceinline__ float4 a0_ld(const float4 *p) {
    float4 r;
    asm("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ float a0_ld(const float *p) {
    float r;
    asm("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(r) : "l"(p));
    return r;
}
template <int VEC> struct A0Lane;
template <> struct A0Lane<4> {
    using T = float4;
    static __device__ __forceinline__ void unpack(const float4 &v, float (&x)[4]) { x[0] = v.x; x[1] = v.y; x[2] = v.z; x[3] = v.w; }
};
template <> struct A0Lane<1> {
    using T = float;
    static __device__ __forceinline__ void unpack(const float &v, float (&x)[1]) { x[0] = v; }
};

struct Axis0K {
    const float *cube;
    int64_t n2, nz;
    int64_t start, end; // frames [start, end)
    int64_t chunk;      // frames per k-chunk
    int ksplit;
    int is_min;
    int64_t r1, r2;     // index offsets of the two arg maps
    int out2_kind;      // 0 none, 1 max/min fold, 2 second arg map (offset r2)
    int sqrt_out;       // stddev
    int use_wcs;
    int wk_table;       // small-image path: per-frame world coordinate tabulated in shared memory
    double crval, crpix, cdelt;
    const float *mask_lo, *mask_hi;
    float *out, *out2;
    void *partials;     // State[ksplit][n2]
    unsigned *flag;     // set to 1 where the reference would panic (mask range outside the cube)
};

// pix2world on one axis: linear WCS evaluated in f64, returned as f32 (fitrs WCS, unit.rs:335-342); without WCS
// the reference falls back to the pixel index itself
__device__ __forceinline__ float a0_world(const Axis0K &a, float pix) {
    if (!a.use_wcs) return pix;
    const double p1 = __dadd_rn((double)pix, __dsub_rn(1.0, a.crpix));
    return (float)__dadd_rn(a.crval, __dmul_rn(a.cdelt, p1));
}

// ---------------------------------------------------------------------------------------------------------------
// functors: State is a POD; init / step (frame k, strictly increasing) / merge (later chunk into earlier) / finish
// ---------------------------------------------------------------------------------------------------------------

// ndarray 0.12.1 var_axis(Axis(0), ddof = 1.0): Welford per pixel in f32, mul_add for m2 (lib.rs:1332-1338)
struct VarSeq {
    struct State { float mean, m2; };
    static constexpr bool kMergeable = false;
    static constexpr bool kNeedsFirst = false;
    static __device__ __forceinline__ void init(State &s, const Axis0K &, int64_t, float) { s.mean = 0.0f; s.m2 = 0.0f; }
    static __device__ __forceinline__ void step(State &s, float x, int64_t k, float wk, const Axis0K &a) {
        (void)wk;
        const float count = (float)(k - a.start + 1); // built by repeated +1.0 in the reference: exact below 2^24
        const float delta = __fsub_rn(x, s.mean);
        s.mean = __fadd_rn(s.mean, __fdiv_rn(delta, count));
        s.m2 = __fmaf_rn(__fsub_rn(x, s.mean), delta, s.m2);
    }
    static __device__ __forceinline__ void merge(State &, const State &, const Axis0K &) {}
    static __device__ __forceinline__ void finish(const State &s, const Axis0K &a, int64_t px) {
        const float dof = __fsub_rn((float)(a.end - a.start), 1.0f);
        float v = __fdiv_rn(s.m2, dof);
        if (a.sqrt_out) v = __fsqrt_rn(v);
        a.out[px] = v;
    }
};

// default mode: f64 sums of d = x - c and d*d with c = the pixel's first frame (shifted data: no cancellation blow-up),
// mergeable by addition; agrees with the two-pass f64 variance to ~1e-12
struct VarComp {
    struct State { double s1, s2; float c; };
    static constexpr bool kMergeable = true;
    static constexpr bool kNeedsFirst = true;
    static __device__ __forceinline__ void init(State &s, const Axis0K &, int64_t, float first) { s.s1 = 0.0; s.s2 = 0.0; s.c = first; }
    static __device__ __forceinline__ void step(State &s, float x, int64_t, float, const Axis0K &) {
        const double d = (double)x - (double)s.c;
        s.s1 += d;
        s.s2 = fma(d, d, s.s2);
    }
    static __device__ __forceinline__ void merge(State &s, const State &o, const Axis0K &) { s.s1 += o.s1; s.s2 += o.s2; }
    static __device__ __forceinline__ void finish(const State &s, const Axis0K &a, int64_t px) {
        const double n = (double)(a.end - a.start);
        double v = (s.s2 - s.s1 * s.s1 / n) / (n - 1.0);
        if (v < 0.0) v = 0.0; // rounding of a constant spectrum
        if (a.sqrt_out) v = sqrt(v);
        a.out[px] = (float)v;
    }
};

// run_create_argmap (lib.rs:1400-1460) and run_minmax (lib.rs:1374-1398) in one pass.
// Arg map: strict comparison from +-Inf, first extreme wins, NaN never wins, nothing won -> 0.0.
// Fold: acc = (acc > y) ? acc : y from -Inf (max) / (acc < y) ? acc : y from +Inf (min): a NaN frame replaces the
// accumulator and the next frame replaces the NaN, so the fold restarts after every NaN.
struct ArgFold {
    struct State { float value; int idx; float fold; int reset; };
    static constexpr bool kMergeable = true;
    static constexpr bool kNeedsFirst = false;
    static __device__ __forceinline__ void init(State &s, const Axis0K &a, int64_t, float) {
        s.value = a.is_min ? __int_as_float(0x7f800000) : __int_as_float(0xff800000);
        s.idx = -1;
        s.fold = s.value;
        s.reset = 0;
    }
    static __device__ __forceinline__ void step(State &s, float x, int64_t k, float, const Axis0K &a) {
        if (a.is_min ? (x < s.value) : (x > s.value)) { s.value = x; s.idx = (int)(k - a.start); }
        if (x != x) s.reset = 1;
        s.fold = (a.is_min ? (s.fold < x) : (s.fold > x)) ? s.fold : x;
    }
    static __device__ __forceinline__ void merge(State &s, const State &o, const Axis0K &a) {
        if (o.idx >= 0 && (a.is_min ? (o.value < s.value) : (o.value > s.value))) { s.value = o.value; s.idx = o.idx; }
        if (o.reset) { s.fold = o.fold; s.reset = 1; }
        else s.fold = (a.is_min ? (s.fold < o.fold) : (s.fold > o.fold)) ? s.fold : o.fold;
    }
    static __device__ __forceinline__ void finish(const State &s, const Axis0K &a, int64_t px) {
        // out = pix2world(trypix, ((k + start) as i64 + range) as f32), or that index itself (lib.rs:1437-1447)
        a.out[px] = s.idx < 0 ? 0.0f : a0_world(a, (float)(a.start + s.idx + a.r1));
        if (a.out2_kind == 1) a.out2[px] = s.fold;
        else if (a.out2_kind == 2) a.out2[px] = s.idx < 0 ? 0.0f : a0_world(a, (float)(a.start + s.idx + a.r2));
    }
};

// run_centroid (lib.rs:1466-1527): flux_sum = sum f, waveimg = sum fl(f * w_k), out = waveimg / flux_sum; all f32,
// sequential from 0.0
struct CentroidSeq {
    struct State { float f, w; };
    static constexpr bool kMergeable = false;
    static constexpr bool kNeedsFirst = false;
    static __device__ __forceinline__ void init(State &s, const Axis0K &, int64_t, float) { s.f = 0.0f; s.w = 0.0f; }
    static __device__ __forceinline__ void step(State &s, float x, int64_t, float wk, const Axis0K &) {
        s.f = __fadd_rn(s.f, x);
        s.w = __fadd_rn(s.w, __fmul_rn(x, wk));
    }
    static __device__ __forceinline__ void merge(State &, const State &, const Axis0K &) {}
    static __device__ __forceinline__ void finish(const State &s, const Axis0K &a, int64_t px) { a.out[px] = __fdiv_rn(s.w, s.f); }
};

// default mode: the product f*w_k is rounded to f32 like the reference, the two sums are accumulated in f64
struct CentroidComp {
    struct State { double f, w; };
    static constexpr bool kMergeable = true;
    static constexpr bool kNeedsFirst = false;
    static __device__ __forceinline__ void init(State &s, const Axis0K &, int64_t, float) { s.f = 0.0; s.w = 0.0; }
    static __device__ __forceinline__ void step(State &s, float x, int64_t, float wk, const Axis0K &) {
        s.f += (double)x;
        s.w += (double)x * (double)wk;
    }
    static __device__ __forceinline__ void merge(State &s, const State &o, const Axis0K &) { s.f += o.f; s.w += o.w; }
    static __device__ __forceinline__ void finish(const State &s, const Axis0K &a, int64_t px) { a.out[px] = (float)(s.w / s.f); }
};

template <class F, int VEC, int U, bool WAVE>
__device__ __forceinline__ void a0_batch(const char *&ptr, size_t stride, int64_t &k, typename F::State (&st)[VEC], const Axis0K &a) {
    using V = typename A0Lane<VEC>::T;
    V x[U];
#pragma unroll
    for (int j = 0; j < U; ++j) x[j] = a0_ld(reinterpret_cast<const V *>(ptr + (size_t)j * stride));
    ptr += (size_t)U * stride;
#pragma unroll
    for (int j = 0; j < U; ++j) {
        float e[VEC];
        A0Lane<VEC>::unpack(x[j], e);
        const float wk = WAVE ? a0_world(a, (float)(k + j)) : 0.0f;
#pragma unroll
        for (int i = 0; i < VEC; ++i) F::step(st[i], e[i], k + j, wk, a);
    }
    k += U;
}

// Order-dependent functors on small images (one pixel per thread, a few warps on the whole GPU): the chain of a pixel cannot
// be split, so a batch of U loads and the U dependent steps that consume it would alternate, each waiting for the other
// (measured 381 us on the MaNGA cube: 1 us of load latency + 0.4 us of Welford steps per 16 frames).  Two register
// buffers keep the next batch in flight while the current one is consumed.
template <class F, int U, bool WAVE>
__device__ __forceinline__ void a0_pipelined(const char *&ptr, size_t stride, int64_t &k, int64_t kend, typename F::State (&st)[1], const Axis0K &a,
                                             const float *wtab) {
    float xa[U], xb[U];
    const char *lp = ptr;
    int64_t kl = k;
    auto load = [&](float (&x)[U]) {
#pragma unroll
        for (int j = 0; j < U; ++j) x[j] = a0_ld(reinterpret_cast<const float *>(lp + (size_t)j * stride));
        lp += (size_t)U * stride;
        kl += U;
    };
    auto consume = [&](const float (&x)[U]) {
#pragma unroll
        for (int j = 0; j < U; ++j) {
            const float wk = WAVE ? (wtab ? wtab[k + j - a.start] : a0_world(a, (float)(k + j))) : 0.0f;
            F::step(st[0], x[j], k + j, wk, a);
        }
        k += U;
    };
    if (kend - kl < U) return;
    load(xa);
    for (;;) {
        const bool more_b = kend - kl >= U;
        if (more_b) load(xb);
        consume(xa);
        if (!more_b) break;
        const bool more_a = kend - kl >= U;
        if (more_a) load(xa);
        consume(xb);
        if (!more_a) break;
    }
    ptr = lp;
}

template <class F, int VEC, int U, bool WAVE, bool PIPE = false>
__global__ void __launch_bounds__(256, 2) axis0_kernel(const Axis0K a) {
    // pix2world of a frame is the same for every pixel and costs three f64 operations: with one pixel per thread (the
    // small-image path) the CTA tabulates it once in shared memory instead of every thread redoing it for every frame
    extern __shared__ float s_wk[];
    const float *wtab = nullptr;
    if constexpr (PIPE && WAVE) {
        if (a.wk_table) {
            for (int64_t i = threadIdx.x; i < a.end - a.start; i += blockDim.x) s_wk[i] = a0_world(a, (float)(a.start + i));
            __syncthreads();
            wtab = s_wk;
        }
    }
    const int64_t px = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * VEC;
    if (px >= a.n2) return;
    const int kc = blockIdx.y;
    int64_t k = a.start + (int64_t)kc * a.chunk;
    const int64_t kend = min(a.end, k + a.chunk);
    typename F::State st[VEC];
    {
        float first[VEC];
#pragma unroll
        for (int i = 0; i < VEC; ++i) first[i] = 0.0f;
        if (F::kNeedsFirst) {
            typename A0Lane<VEC>::T v = a0_ld(reinterpret_cast<const typename A0Lane<VEC>::T *>(a.cube + a.start * a.n2 + px));
            A0Lane<VEC>::unpack(v, first);
        }
#pragma unroll
        for (int i = 0; i < VEC; ++i) F::init(st[i], a, px + i, first[i]);
    }
    const size_t stride = (size_t)a.n2 * sizeof(float);
    const char *ptr = reinterpret_cast<const char *>(a.cube + px) + (size_t)k * stride;
    if constexpr (PIPE && VEC == 1) a0_pipelined<F, 2 * U, WAVE>(ptr, stride, k, kend, st, a, wtab);
    while (kend - k >= U) a0_batch<F, VEC, U, WAVE>(ptr, stride, k, st, a);
    if (U > 4 && kend - k >= 4) a0_batch<F, VEC, 4, WAVE>(ptr, stride, k, st, a);
    if (U > 2 && kend - k >= 2) a0_batch<F, VEC, 2, WAVE>(ptr, stride, k, st, a);
    while (kend - k >= 1) a0_batch<F, VEC, 1, WAVE>(ptr, stride, k, st, a);
    if (a.ksplit == 1) {
#pragma unroll
        for (int i = 0; i < VEC; ++i) F::finish(st[i], a, px + i);
    } else {
        typename F::State *dst = static_cast<typename F::State *>(a.partials) + (size_t)kc * a.n2 + px;
#pragma unroll
        for (int i = 0; i < VEC; ++i) dst[i] = st[i];
    }
}

template <class F> __global__ void __launch_bounds__(256) axis0_merge_kernel(const Axis0K a) {
    const int64_t px = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (px >= a.n2) return;
    const typename F::State *src = static_cast<const typename F::State *>(a.partials) + px;
    typename F::State st = src[0];
    for (int c = 1; c < a.ksplit; ++c) F::merge(st, src[(size_t)c * a.n2], a);
    F::finish(st, a, px);
}

// run_centroid_with_mask (lib.rs:1529-1582): per-pixel range [start_mask as usize, end_mask as usize), sequential f32
// sums, wavelength from pix2world(2, k).  slice_axis panics when start > end or end > nz: flagged, never read.
template <bool COMP> __global__ void __launch_bounds__(256) centroid_mask_kernel(const Axis0K a) {
    const int64_t px = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (px >= a.n2) return;
    auto as_usize = [](float f) -> int64_t { // Rust `as usize`: truncate, saturate, NaN -> 0
        if (!(f > 0.0f)) return 0;
        if (f >= 9.2233720368547758e18f) return INT64_MAX;
        return (int64_t)f;
    };
    const int64_t lo = as_usize(a.mask_lo[px]), hi = as_usize(a.mask_hi[px]);
    if (lo > hi || hi > a.nz) {
        atomicMax(a.flag, 1u);
        a.out[px] = 0.0f;
        return;
    }
    const float *p = a.cube + px;
    if (COMP) {
        double f = 0.0, w = 0.0;
        for (int64_t k = lo; k < hi; ++k) {
            const float x = a0_ld(p + k * a.n2);
            f += (double)x;
            w += (double)x * (double)a0_world(a, (float)k);
        }
        a.out[px] = (float)(w / f);
    } else {
        float f = 0.0f, w = 0.0f;
        for (int64_t k = lo; k < hi; ++k) {
            const float x = a0_ld(p + k * a.n2);
            f = __fadd_rn(f, x);
            w = __fadd_rn(w, __fmul_rn(x, a0_world(a, (float)k)));
        }
        a.out[px] = __fdiv_rn(w, f);
    }
}

// ---------------------------------------------------------------------------------------------------------------
// run_median (lib.rs:1340-1372): per-pixel median of n = end - start frames; n odd -> middle element, n even ->
// fl((v[n/2] + v[n/2-1]) / 2).  The reference sorts with partial_cmp().unwrap(): it panics if any value is NaN
// (and n > 1).  Values are compared as order-preserving integer keys, so the result is the reference's bit for bit
// (among equal values -0.0 sorts before +0.0, which the sum of the two middles cannot tell apart).
//
// All variants read the sub-cube from HBM exactly once, staged into shared memory with cp.async (thousands of copies in
// flight per CTA, no registers held):
//   n <= 128 median_sort_kernel: lane = spaxel, the column is loaded into registers (coalesced) and sorted by an unrolled
//            bitonic network.
//   larger   median_select_kernel: a CTA stages P spaxels transposed to [P][npad] (npad = n padded so that both the
//            staging stores and the per-spectrum loads are bank-conflict free); one warp per spaxel runs a radix select
//            over key - min(key): the digits cover only the range the spectrum really spans (typically 3 passes of 8
//            bits, well spread histogram bins).
//   long windows on large images, and any window beyond shared memory: median_stream_kernel (below) streams the cube twice
//            with coalesced loads instead of staging it.
// ---------------------------------------------------------------------------------------------------------------
struct MedianK {
    const float *cube;
    int64_t n2;
    int64_t start, n;
    int P;          // spaxels per CTA
    int npad;       // words per staged spectrum (select kernel)
    float *out;
    unsigned *flag; // 1: NaN present (reference panics)
    unsigned long long *trace; // AFK_MEDIAN_TRACE=1 (diagnostics): clock cycles per phase summed over the CTAs, else nullptr
};

__device__ __forceinline__ void ms_trace(const MedianK &a, int phase, long long &t) {
    if (a.trace && threadIdx.x == 0) {
        const long long now = clock64();
        atomicAdd(a.trace + phase, (unsigned long long)(now - t));
        t = now;
    }
}

__device__ __forceinline__ unsigned med_key(float f) {
    const unsigned b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float med_unkey(unsigned k) {
    const unsigned b = (k & 0x80000000u) ? (k & 0x7FFFFFFFu) : ~k;
    return __uint_as_float(b);
}
__device__ __forceinline__ float med_finish(unsigned klo, unsigned khi, bool even) {
    return even ? __fdiv_rn(__fadd_rn(med_unkey(khi), med_unkey(klo)), 2.0f) : med_unkey(klo);
}

constexpr int kMedSortPixels = 128; // spaxels (= threads) per CTA of the sorting-network kernel
constexpr int kMedRankMaxN = 128;

// Windows of at most 128 frames (the template-sized case): lane = spaxel, the whole column lives in registers -- coalesced
// scalar loads straight from global memory (a warp reads 128 contiguous bytes of a frame, every frame of the window in
// flight at once), padded with the largest key to the next power of two N, sorted by a fully unrolled bitonic network
// (N/2 log2 N (log2 N + 1) / 2 compare-exchanges of two instructions each: 672 for N = 64, against 64^2 x 5 for ranking
// by counting out of shared memory, the first version: 0.51 -> see profiles/r1_axis0_ops.json).
template <int N>
__global__ void __launch_bounds__(kMedSortPixels) median_sort_kernel(const MedianK a) {
    const int64_t px = (int64_t)blockIdx.x * kMedSortPixels + threadIdx.x;
    if (px >= a.n2) return;
    const int n = (int)a.n;
    const float *col = a.cube + a.start * a.n2 + px;
    unsigned key[N];
#pragma unroll
    for (int k = 0; k < N; ++k) key[k] = k < n ? __float_as_uint(__ldg(col + (size_t)k * a.n2)) : 0x7FFFFFFFu; // pads map to 0xFFFFFFFF
    bool nan = false;
#pragma unroll
    for (int k = 0; k < N; ++k) {
        const float f = __uint_as_float(key[k]);
        nan |= (k < n) && (f != f);
        key[k] = med_key(f);
    }
    if (nan && n > 1) atomicMax(a.flag, 1u);
#pragma unroll
    for (int size = 2; size <= N; size <<= 1) {
#pragma unroll
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
#pragma unroll
            for (int i = 0; i < N; ++i) {
                const int j = i ^ stride;
                if (j > i) {
                    const bool up = (i & size) == 0;
                    const unsigned lo = min(key[i], key[j]), hi = max(key[i], key[j]);
                    key[i] = up ? lo : hi;
                    key[j] = up ? hi : lo;
                }
            }
        }
    }
    // the two middles: compile-time register indices, selected by comparison (a dynamic index would spill the array)
    const int r_lo = (n - 1) / 2, r_hi = n / 2;
    unsigned klo = 0u, khi = 0u;
#pragma unroll
    for (int k = 0; k < N; ++k) {
        if (k == r_lo) klo = key[k];
        if (k == r_hi) khi = key[k];
    }
    a.out[px] = med_finish(klo, khi, (n & 1) == 0);
}

// One radix-select digit step: given the histogram of the candidates' current digit (256 bins), find the bin holding
// rank kth (0-based among the candidates), make kth relative to that bin, return the bin and its population.
__device__ __forceinline__ unsigned med_pick_digit(const unsigned *hist, unsigned &kth, unsigned &bin_count, int lane) {
    // lane l owns bins [8l, 8l+8); exclusive scan of the lane totals
    unsigned c[8], tot = 0u;
#pragma unroll
    for (int j = 0; j < 8; ++j) { c[j] = hist[lane * 8 + j]; tot += c[j]; }
    unsigned incl = tot;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    unsigned excl = incl - tot;
    int digit = -1;
    unsigned below = 0u, cnt = 0u;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        if (digit < 0 && kth < excl + c[j]) { digit = lane * 8 + j; below = excl; cnt = c[j]; }
        excl += c[j];
    }
    // cumulative counts are monotonic: the lowest lane that found a bin has the first bin crossing kth
    const unsigned ballot = __ballot_sync(0xffffffffu, digit >= 0);
    const int src = __ffs(ballot) - 1;
    digit = __shfl_sync(0xffffffffu, digit, src);
    below = __shfl_sync(0xffffffffu, below, src);
    bin_count = __shfl_sync(0xffffffffu, cnt, src);
    kth -= below;
    return (unsigned)digit;
}

// A spectrum is solved by a group of G warps (G = 8 / min(P, 8): with few spaxels per CTA all eight warps still work);
// the group synchronises on its own named barrier, or with __syncwarp when it is a single warp.
__device__ __forceinline__ void med_group_sync(int G, int grp) {
    if (G == 1) __syncwarp();
    else asm volatile("bar.sync %0, %1;" ::"r"(1 + grp), "r"(G * 32) : "memory");
}

// keys: the n order-preserving keys of one spectrum, contiguous; digits are taken from key - kmin, span = kmax - kmin.
// Finds the offset keys of ranks r and (when want_next) r + 1; every thread of the group returns the same values.
// hist = 256 words, red = 8 words of shared memory owned by the group.
__device__ void med_select_group(const unsigned *keys, int n, unsigned kmin, unsigned span, unsigned r, bool want_next, unsigned *hist,
                                 unsigned *red, int gt, int gsize, int lane, int G, int grp, bool vec4, unsigned &k_r, unsigned &k_next) {
    int bits_left = 32 - __clz(span); // 0 when every key is equal
    unsigned prefix = 0u, kth = r, eq_count = (unsigned)n;
    while (bits_left > 0) {
        const int w = bits_left < 8 ? bits_left : 8;
        const int shift = bits_left - w;
        const unsigned dmask = (1u << w) - 1u;
        const unsigned cmask = bits_left >= 32 ? 0u : ~((1u << bits_left) - 1u);
        for (int b = gt; b < 256; b += gsize) hist[b] = 0u;
        med_group_sync(G, grp);
        if (vec4) {
            // 16-byte shared loads: a quarter of the load instructions for the same bytes (the select is bound by the
            // shared-memory instruction queue, not by its bandwidth); two loads in flight per thread
            const uint4 *k4 = reinterpret_cast<const uint4 *>(keys);
            const int n4 = (n + 3) >> 2;
            for (int base = 0; base < n4; base += 2 * gsize) {
                const int q0 = base + gt, q1 = q0 + gsize;
                uint4 v0 = make_uint4(0, 0, 0, 0), v1 = v0;
                if (q0 < n4) v0 = k4[q0];
                if (q1 < n4) v1 = k4[q1];
                const unsigned kk[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    const int i = 4 * (u < 4 ? q0 : q1) + (u & 3);
                    const unsigned key = kk[u] - kmin;
                    if (i < n && (key & cmask) == prefix) atomicAdd(&hist[(key >> shift) & dmask], 1u);
                }
            }
        } else {
            for (int base = 0; base < n; base += 4 * gsize) {
                unsigned key[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int i = base + u * gsize + gt;
                    key[u] = i < n ? keys[i] - kmin : 0u;
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int i = base + u * gsize + gt;
                    if (i < n && (key[u] & cmask) == prefix) atomicAdd(&hist[(key[u] >> shift) & dmask], 1u);
                }
            }
        }
        med_group_sync(G, grp);
        const unsigned digit = med_pick_digit(hist, kth, eq_count, lane); // every warp of the group reads the same histogram
        prefix |= digit << shift;
        bits_left = shift;
        med_group_sync(G, grp); // nobody may still be reading when the histogram is cleared again
    }
    k_r = prefix;
    // eq_count keys equal k_r and kth is r's position among them
    if (!want_next || kth + 1 < eq_count) { k_next = prefix; return; } // group-uniform
    unsigned m = 0xFFFFFFFFu; // smallest offset key above k_r
    for (int i = gt; i < n; i += gsize) {
        const unsigned key = keys[i] - kmin;
        if (key > prefix && key < m) m = key;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = min(m, __shfl_xor_sync(0xffffffffu, m, o));
    if (G > 1) {
        if (lane == 0) red[gt >> 5] = m;
        med_group_sync(G, grp);
        m = 0xFFFFFFFFu;
        for (int w = 0; w < G; ++w) m = min(m, red[w]);
        med_group_sync(G, grp);
    }
    k_next = m;
}

__global__ void __launch_bounds__(256) median_select_kernel(const MedianK a) {
    extern __shared__ __align__(16) unsigned s_keys[]; // [P][npad] staged keys | 8 x 256 histogram words | 8 x 16 reduction words
    const int P = a.P, npad = a.npad, n = (int)a.n;
    const int64_t px0 = (int64_t)blockIdx.x * P;
    unsigned *s_hist = s_keys + (size_t)P * npad;
    unsigned *s_red = s_hist + 8 * 256;
    const int tid = threadIdx.x;
    // P is a power of two: frame k = i >> log2(P), spaxel j = i & (P - 1); no divisions in the staging loop
    const int logp = 31 - __clz(P);
    const unsigned total = (unsigned)n * (unsigned)P; // < 2^16 words of shared memory
    const float *src0 = a.cube + a.start * a.n2 + px0;
    const unsigned dst0 = (unsigned)__cvta_generic_to_shared(s_keys);
    for (unsigned i = tid; i < total; i += 256) {
        const unsigned k = i >> logp, j = i & (unsigned)(P - 1);
        if (px0 + j < a.n2) {
            const float *src = src0 + (size_t)k * a.n2 + j;
            const unsigned dst = dst0 + 4u * (j * (unsigned)npad + k);
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src) : "memory");
        }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
    const int warp = tid >> 5, lane = tid & 31;
    const int G = P >= 8 ? 1 : 8 / P, ngroups = 8 / G;
    const int grp = warp / G, gt = (warp % G) * 32 + lane, gsize = G * 32;
    unsigned *hist = s_hist + grp * 256, *red = s_red + grp * 16;
    bool nan = false;
    for (int j = grp; j < P; j += ngroups) {
        const int64_t px = px0 + j;
        if (px >= a.n2) continue; // group-uniform
        unsigned *keys = s_keys + (size_t)j * npad;
        unsigned kmin = 0xFFFFFFFFu, kmax = 0u;
        const bool vec4 = (npad & 3) == 0; // every staged spectrum then starts on a 16-byte boundary
        if (vec4) {
            uint4 *k4 = reinterpret_cast<uint4 *>(keys);
            const int n4 = (n + 3) >> 2;
            for (int q = gt; q < n4; q += gsize) {
                const uint4 v = k4[q];
                unsigned kk[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const float f = __uint_as_float(kk[u]);
                    kk[u] = med_key(f);
                    if (4 * q + u < n) { // the words between n and npad are never staged
                        nan |= (f != f);
                        kmin = min(kmin, kk[u]);
                        kmax = max(kmax, kk[u]);
                    }
                }
                k4[q] = make_uint4(kk[0], kk[1], kk[2], kk[3]);
            }
        } else {
            for (int i = gt; i < n; i += gsize) {
                const float f = __uint_as_float(keys[i]);
                nan |= (f != f);
                const unsigned key = med_key(f);
                keys[i] = key;
                kmin = min(kmin, key);
                kmax = max(kmax, key);
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            kmin = min(kmin, __shfl_xor_sync(0xffffffffu, kmin, o));
            kmax = max(kmax, __shfl_xor_sync(0xffffffffu, kmax, o));
        }
        if (G > 1) {
            if (lane == 0) { red[2 * (gt >> 5)] = kmin; red[2 * (gt >> 5) + 1] = kmax; }
            med_group_sync(G, grp);
            for (int w = 0; w < G; ++w) { kmin = min(kmin, red[2 * w]); kmax = max(kmax, red[2 * w + 1]); }
        }
        med_group_sync(G, grp); // the converted keys are visible to the whole group; `red` may be reused
        unsigned klo, khi;
        med_select_group(keys, n, kmin, kmax - kmin, (unsigned)((n - 1) / 2), (n & 1) == 0, hist, red, gt, gsize, lane, G, grp, vec4, klo, khi);
        if (gt == 0) a.out[px] = med_finish(klo + kmin, khi + kmin, (n & 1) == 0);
    }
    if (nan && n > 1) atomicMax(a.flag, 1u); // a single frame is never compared by the sort
}

// ---------------------------------------------------------------------------------------------------------------
// Warp-synchronous select on a short list in shared memory (the last phase of median_stream_kernel) -- no atomics, no
// barriers.  Shared-memory atomics run at about one per clock per SM (the histogram select above spends a clock per
// voxel and pass on them); compare-and-count in registers is an order of magnitude cheaper per key when the pivots are
// good, so they come from the data:
//   round: 32 jittered stratified samples, one per lane, bitonic-sorted across the warp; 7 of them around the sample
//          median are the pivots.  Count pass: c_j = #(key < p_j) in registers.  The bin (lo, hi) that holds rank r is then
//          compacted IN PLACE (ballot / popc offsets; every load of a chunk precedes its stores, which land in the
//          already-read prefix), together with the number of keys equal to lo (they are dropped: a bin made of
//          duplicates still shrinks) and the smallest key >= hi (the even-n successor).
//   Rounds repeat on the compacted list (~240 -> ~15 keys) until it fits one key per lane, then it is ranked by counting.
//   A pivot set that misses the median only makes the next list longer: exactness never depends on the data.
// ---------------------------------------------------------------------------------------------------------------
constexpr int kMedPivots = 7;
constexpr unsigned kFull = 0xffffffffu;

// ascending bitonic sort of one value per lane
__device__ __forceinline__ unsigned med_warp_sort(unsigned v, int lane) {
#pragma unroll
    for (int k = 2; k <= 32; k <<= 1) {
#pragma unroll
        for (int j = k >> 1; j > 0; j >>= 1) {
            const unsigned o = __shfl_xor_sync(kFull, v, j);
            const bool up = ((lane & k) == 0) == ((lane & j) == 0); // keep the smaller one
            v = up ? min(v, o) : max(v, o);
        }
    }
    return v;
}

struct MedState {
    int m;            // keys in the list
    unsigned r;       // wanted rank within the list
    unsigned above;   // smallest key known to sort after every key of the list
    bool has_above;
};

// One round on list `keys[0, st.m)`.  Returns true when the answer is known (k_r / k_next set).
__device__ __forceinline__ bool med_round(unsigned *keys, MedState &st, bool want_next, int lane, unsigned seed, unsigned &k_r, unsigned &k_next) {
    const int m = st.m;
    // ---- pivots: stratified sample, one key per lane, jittered so that periodic spectra cannot alias the stride ----
    const int stride = m >> 5; // m > 32
    unsigned h = (seed ^ (unsigned)lane * 0x9E3779B9u) * 0x85EBCA6Bu;
    h ^= h >> 15;
    const int pos = (int)(((long long)lane * m) >> 5) + (int)(h % (unsigned)stride);
    const unsigned smp = med_warp_sort(keys[pos], lane);
    // 7 pivots two sample ranks apart, centred on where the wanted rank is expected among the 32 sorted samples (the sample
    // rank of list rank r has sigma <= 2.8): the bin that holds r is then ~1/16 of the list wherever r lies, not only when
    // it is near the middle
    int centre = (int)(((long long)st.r * 32 + m / 2) / m);
    centre = centre < 6 ? 6 : (centre > 25 ? 25 : centre);
    unsigned pv[kMedPivots];
#pragma unroll
    for (int j = 0; j < kMedPivots; ++j) pv[j] = __shfl_sync(kFull, smp, centre - 6 + 2 * j);
    // ---- count pass ----
    unsigned c[kMedPivots];
#pragma unroll
    for (int j = 0; j < kMedPivots; ++j) c[j] = 0u;
    const bool vec = (m >= 256) && ((reinterpret_cast<uintptr_t>(keys) & 15u) == 0);
    const int nvec = vec ? (m & ~255) : 0; // keys handled 8 per lane
    for (int base = 0; base < nvec; base += 256) {
        const uint4 *k4 = reinterpret_cast<const uint4 *>(keys + base);
        const uint4 v0 = k4[lane], v1 = k4[lane + 32];
        const unsigned kk[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
#pragma unroll
        for (int u = 0; u < 8; ++u)
#pragma unroll
            for (int j = 0; j < kMedPivots; ++j) c[j] += kk[u] < pv[j];
    }
    for (int i = nvec + lane; i < m; i += 32) {
        const unsigned key = keys[i];
#pragma unroll
        for (int j = 0; j < kMedPivots; ++j) c[j] += key < pv[j];
    }
#pragma unroll
    for (int j = 0; j < kMedPivots; ++j) c[j] = __reduce_add_sync(kFull, c[j]);
    // ---- bin of rank r: b = number of pivots with c_j <= r ----

    unsigned below = 0u, lo = 0u, hi = 0u;
    bool has_lo = false, has_hi = false;
#pragma unroll
    for (int j = 0; j < kMedPivots; ++j) {
        if (c[j] <= st.r) { below = c[j]; lo = pv[j]; has_lo = true; }
    }
#pragma unroll
    for (int j = kMedPivots - 1; j >= 0; --j) {
        if (c[j] > st.r) { hi = pv[j]; has_hi = true; }
    }

    // ---- compaction in place: keep lo < key < hi; count key == lo; min of key >= hi ----
    unsigned n_eq = 0u, hi_min = kFull, out = 0u;
    bool hi_any = false;
    __syncwarp();
    for (int base = 0; base < m; base += 256) {
        unsigned kk[8];
        unsigned valid = 0u;
        if (base < nvec) {
            const uint4 *k4 = reinterpret_cast<const uint4 *>(keys + base);
            const uint4 v0 = k4[lane], v1 = k4[lane + 32];
            kk[0] = v0.x; kk[1] = v0.y; kk[2] = v0.z; kk[3] = v0.w; kk[4] = v1.x; kk[5] = v1.y; kk[6] = v1.z; kk[7] = v1.w;
            valid = 0xFFu;
        } else {
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const int i = base + u * 32 + lane;
                kk[u] = 0u;
                if (i < m) { kk[u] = keys[i]; valid |= 1u << u; }
            }
        }
        __syncwarp(); // every load of this chunk precedes its stores
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const bool ok = (valid >> u) & 1u;
            const unsigned key = kk[u];
            const bool ge_hi = ok && has_hi && key >= hi;
            const bool le_lo = ok && has_lo && key <= lo;
            if (ok && has_lo && key == lo) ++n_eq;
            if (ge_hi) { hi_min = min(hi_min, key); hi_any = true; }
            const bool keep = ok && !ge_hi && !le_lo;
            const unsigned bal = __ballot_sync(kFull, keep);
            if (keep) keys[out + __popc(bal & ((1u << lane) - 1u))] = key;
            out += __popc(bal);
        }
    }
    __syncwarp();
    n_eq = __reduce_add_sync(kFull, n_eq);
    if (want_next) {
        hi_min = __reduce_min_sync(kFull, hi_min);
        if (__any_sync(kFull, hi_any)) { st.above = hi_min; st.has_above = true; } // keys >= hi sort below every earlier `above`
    }
    unsigned r = st.r - below;
    if (r < n_eq) { // the wanted key is lo itself
        k_r = lo;
        if (!want_next || r + 1 < n_eq) { k_next = lo; return true; }
        // successor: smallest kept key, else `above`
        unsigned mn = kFull;
        for (int i = lane; i < (int)out; i += 32) mn = min(mn, keys[i]);
        mn = __reduce_min_sync(kFull, mn);
        k_next = out > 0u ? mn : st.above;
        return true;
    }
    st.r = r - n_eq;
    st.m = (int)out;
    return false;
}

// list of at most 32 keys: rank by counting
__device__ __forceinline__ void med_rank32(const unsigned *keys, const MedState &st, bool want_next, int lane, unsigned &k_r, unsigned &k_next) {
    const int m = st.m;
    const unsigned mine = lane < m ? keys[lane] : kFull;
    unsigned lt = 0u, le = 0u;
    for (int j = 0; j < m; ++j) {
        const unsigned o = __shfl_sync(kFull, mine, j);
        lt += o < mine;
        le += o <= mine;
    }
    const unsigned b0 = __ballot_sync(kFull, lane < m && lt <= st.r && st.r < le);
    k_r = __shfl_sync(kFull, mine, __ffs(b0) - 1);
    k_next = k_r;
    if (!want_next) return;
    if (st.r + 1 < (unsigned)m) {
        const unsigned b1 = __ballot_sync(kFull, lane < m && lt <= st.r + 1 && st.r + 1 < le);
        k_next = __shfl_sync(kFull, mine, __ffs(b1) - 1);
    } else {
        k_next = st.above;
    }
}

// ---------------------------------------------------------------------------------------------------------------
// median_stream_kernel: windows longer than 128 frames, any length, any alignment.
//
// Staging whole spectra in shared memory limits a CTA to a handful of spaxels, i.e. to 16-32 contiguous bytes per frame:
// measured 0.75 TB/s for the staging alone on the ALMA cube, whatever the select does afterwards (the memory system
// is bound by the number of sector requests, not by their bytes).  Here the cube is STREAMED like the other axis-0
// reductions -- lane = spaxel, a warp reads 128 contiguous bytes of a frame, 16 frames in flight per thread -- and
// only a small candidate list per spaxel ever lives in shared memory:
//   A  32 stratified sample frames per spaxel -> sorted by a warp -> 7 pivots around the sample median (registers)
//   B  stream the window: c_j = #(key < pivot_j) per spaxel (compare + add in registers, no atomics)
//      -> the bin (lo, hi) that holds the wanted rank; about 1/16 of the keys fall into it
//   C  stream the window again (largely from L2): append the keys of the bin to the spaxel's list in shared memory (one
//      shared atomic per APPENDED key), count the keys equal to lo, track the smallest key >= hi (even-n successor)
//   D  one warp per spaxel finishes on the list with the warp-synchronous rounds of med_round / med_rank32
// A bin that overflows the list (heavy ties, unlucky pivots) is refined by its warp against the global column with pivots
// sampled from the bin itself, until it fits: exactness never depends on the data.
// ---------------------------------------------------------------------------------------------------------------
constexpr int kMsPix = 32;            // spaxels per CTA (= lanes)
constexpr int kMsWarps = 16;          // 2 CTAs x 512 threads x 16 loads in flight = 64 KB per SM (one 1024-thread CTA per SM, whose tile would stay in L2, measured slower: 12.8 vs 11.5 ms)
constexpr int kMsCap = 800;           // candidate slots per spaxel (20 % of an ALMA window: a 4-rank bin of 64 samples overflows 4e-4 of the time)
constexpr int kMsRow = kMsCap + 1;    // odd row stride: neighbouring spaxels start in different banks
constexpr int kMsU = 16;              // frames in flight per thread
constexpr int kMsUC = 16;             // ... in the collect pass: 20 -> 10.9 ms, 32 -> 15.3 ms (spills under the 64-register cap); 16 -> 9.8 ms
constexpr int64_t kMedStreamWideBins = 4200;         // up to here a 4-rank bin (6 % of the window) fits the list three times over
constexpr int64_t kMedStreamMinFrames = 129;         // every window beyond the sorting network, on images large enough (measured 100..3000 frames: profiles/r1_median_phases.txt)
constexpr int64_t kMedStreamMinPixels = 2 * 148 * 32; // fewer spaxels: not enough 32-spaxel CTAs for two per SM

__device__ __forceinline__ unsigned ms_ld(const float *p) {
    unsigned r;
    asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(r) : "l"(p));
    return r;
}
// L2 eviction priorities (a 32-bit load takes them through a cache-hint policy): the count pass marks its lines
// evict_last -- the collect pass reads them again -- and the collect pass evict_first: last use.
__device__ __forceinline__ uint64_t ms_policy_keep() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ uint64_t ms_policy_last() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ unsigned ms_ld_hint(const float *p, uint64_t pol) {
    unsigned r;
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.u32 %0, [%1], %2;" : "=r"(r) : "l"(p), "l"(pol));
    return r;
}
__device__ __forceinline__ unsigned ms_key(unsigned b) { return b ^ ((unsigned)((int)b >> 31) | 0x80000000u); } // == med_key, 2 instructions

// c += (key < pivot) as compare + predicated add: two instructions (the C form compiles to three)
__device__ __forceinline__ void ms_count(unsigned &c, unsigned key, unsigned pivot) {
    asm("{\n\t.reg .pred p;\n\tsetp.lt.u32 p, %1, %2;\n\t@p add.u32 %0, %0, 1;\n\t}" : "+r"(c) : "r"(key), "r"(pivot));
}

struct MsBin { // keys strictly inside (lo, hi) are candidates; sentinels 0 / 0xFFFFFFFF only collide with NaN keys
    unsigned lo, hi, below;
    bool has_lo;
};

// bin of rank r from cumulative counts c_j = #(key < pv_j) (pv ascending)
template <int NP> __device__ __forceinline__ MsBin ms_pick_bin(const unsigned (&c)[NP], const unsigned (&pv)[NP], unsigned r) {
    MsBin b{0u, kFull, 0u, false};
#pragma unroll
    for (int j = 0; j < NP; ++j)
        if (c[j] <= r) { b.lo = pv[j]; b.below = c[j]; b.has_lo = true; }
#pragma unroll
    for (int j = NP - 1; j >= 0; --j)
        if (c[j] > r) b.hi = pv[j];
    return b;
}

// One refinement round of an overflowing bin by a single warp, against the global column (lane = frame).
// On entry the list holds kMsCap keys of the bin (lo, hi) and r is the rank among the keys strictly inside it.
// Returns true when the answer is known.
__device__ bool ms_refine(const float *col, int64_t n2, int n, unsigned *list, unsigned &lo, unsigned &hi, bool &has_lo, unsigned &r,
                          unsigned &m, unsigned &above, bool want_next, int lane, unsigned seed, unsigned &k_r, unsigned &k_next) {
    unsigned h = (seed ^ (unsigned)lane * 0x9E3779B9u) * 0x85EBCA6Bu;
    h ^= h >> 15;
    const unsigned smp = med_warp_sort(list[lane * (kMsCap / 32) + (h % (kMsCap / 32))], lane);
    unsigned pv[kMedPivots], c[kMedPivots];
#pragma unroll
    for (int j = 0; j < kMedPivots; ++j) { pv[j] = __shfl_sync(kFull, smp, 10 + 2 * j); c[j] = 0u; }
    for (int k = lane; k < n; k += 32) {
        const unsigned key = ms_key(ms_ld(col + (size_t)k * n2));
        if (key > lo && key < hi) {
#pragma unroll
            for (int j = 0; j < kMedPivots; ++j) c[j] += key < pv[j];
        }
    }
#pragma unroll
    for (int j = 0; j < kMedPivots; ++j) c[j] = __reduce_add_sync(kFull, c[j]);
    const MsBin b = ms_pick_bin<kMedPivots>(c, pv, r);
    const unsigned nlo = b.has_lo ? b.lo : lo, nhi = b.hi < hi ? b.hi : hi;
    unsigned n_eq = 0u, hi_min = kFull, out = 0u;
    __syncwarp();
    for (int k0 = 0; k0 < n; k0 += 32) {
        const int k = k0 + lane;
        const unsigned key = k < n ? ms_key(ms_ld(col + (size_t)k * n2)) : 0u;
        const bool in_old = k < n && key > lo && key < hi;
        if (in_old && b.has_lo && key == nlo) ++n_eq;
        if (in_old && key >= nhi) hi_min = min(hi_min, key);
        const bool keep = in_old && key > nlo && key < nhi;
        const unsigned bal = __ballot_sync(kFull, keep);
        const unsigned at = out + __popc(bal & ((1u << lane) - 1u));
        if (keep && at < (unsigned)kMsCap) list[at] = key;
        out += __popc(bal);
    }
    __syncwarp();
    n_eq = __reduce_add_sync(kFull, n_eq);
    hi_min = __reduce_min_sync(kFull, hi_min);
    if (hi_min != kFull) above = hi_min;
    const unsigned rr = r - b.below;
    if (b.has_lo && rr < n_eq) {
        k_r = nlo;
        if (!want_next || rr + 1 < n_eq) { k_next = nlo; return true; }
        unsigned mn = kFull; // successor of nlo inside the old bin, else `above`
        for (int k = lane; k < n; k += 32) {
            const unsigned key = ms_key(ms_ld(col + (size_t)k * n2));
            if (key > nlo && key < hi) mn = min(mn, key);
        }
        mn = __reduce_min_sync(kFull, mn);
        k_next = mn != kFull ? mn : above;
        return true;
    }
    r = rr - (b.has_lo ? n_eq : 0u);
    lo = nlo; hi = nhi; has_lo = has_lo || b.has_lo;
    m = out;
    return false;
}

// NP pivots, G sample ranks apart, centred on the sample median: <7, 4> for windows up to ~4200 frames (6 % bins), <15, 2>
// beyond (3 % bins: a bin must fit the 768-slot list with margin; the pivots must still span +-3 sigma of the median's rank)
template <int NP, int G>
__global__ void __launch_bounds__(kMsPix * kMsWarps, 2) median_stream_kernel(const MedianK a) {
    extern __shared__ unsigned s_list[]; // [kMsPix][kMsRow] candidate lists; phases A/B alias samples and per-warp counts onto it
    __shared__ unsigned s_piv[kMsPix][NP + 1];
    __shared__ unsigned s_lo[kMsPix], s_hi[kMsPix], s_below[kMsPix], s_haslo[kMsPix], s_cnt[kMsPix];
    __shared__ unsigned s_eq[kMsWarps][kMsPix], s_himin[kMsWarps][kMsPix];
    const int n = (int)a.n;
    const int tid = threadIdx.x, g = tid >> 5, lane = tid & 31;
    const int64_t px0 = (int64_t)blockIdx.x * kMsPix;
    const int64_t mypx = px0 + lane;
    const bool active = mypx < a.n2;
    const float *col = a.cube + a.start * a.n2 + (active ? mypx : a.n2 - 1); // inactive lanes read a valid address
    const unsigned r0 = (unsigned)((n - 1) / 2);
    const bool even = (n & 1) == 0;
    bool nan = false;
    unsigned lo = 0u, hi = kFull;
    if (tid < kMsPix) { s_cnt[tid] = 0u; s_lo[tid] = 0u; s_hi[tid] = kFull; s_below[tid] = 0u; s_haslo[tid] = 0u; }
    if (n > kMsCap) {
        // ---- A: 64 stratified sample frames per spaxel (the same frames for the whole CTA: coalesced) -> pivots ----
        unsigned *s_smp = s_list; // [kMsPix][65]
        const int stride = n >> 6; // n > kMsCap >= 64
        for (int i = g; i < 64; i += kMsWarps) {
            unsigned h = ((unsigned)blockIdx.x * 0x9E3779B9u + (unsigned)i * 0x85EBCA6Bu) * 0xC2B2AE35u;
            h ^= h >> 16;
            const int f = (int)(((long long)i * n) >> 6) + (int)(h % (unsigned)stride);
            s_smp[lane * 65 + i] = (ms_key(ms_ld(col + (size_t)f * a.n2)) & ~63u) | (unsigned)i;
        }
        __syncthreads();
        // a warp ranks the 64 samples of a spaxel by counting.  The low 6 bits of each sample were replaced by its index
        // when it was stored, so the 64 values are distinct and one compare ranks them; a pivot does not have to be a key
        // of the spectrum.  The pivots are the samples of ranks 32 + (j - (NP-1)/2) G.
        for (int p = g; p < kMsPix; p += kMsWarps) {
            const unsigned *sp = s_smp + p * 65;
            const unsigned v0 = sp[lane], v1 = sp[lane + 32];
            unsigned r0s = 0u, r1s = 0u;
#pragma unroll 8
            for (int i = 0; i < 64; ++i) {
                const unsigned o = sp[i];
                ms_count(r0s, o, v0);
                ms_count(r1s, o, v1);
            }
            constexpr unsigned kFirst = 32u - (unsigned)((NP - 1) / 2 * G), kLast = kFirst + (unsigned)((NP - 1) * G);
            if (r0s >= kFirst && r0s <= kLast && (r0s - kFirst) % G == 0u) s_piv[p][(r0s - kFirst) / G] = v0;
            if (r1s >= kFirst && r1s <= kLast && (r1s - kFirst) % G == 0u) s_piv[p][(r1s - kFirst) / G] = v1;
        }
        __syncthreads();
        unsigned pv[NP], c[NP];
#pragma unroll
        for (int j = 0; j < NP; ++j) { pv[j] = s_piv[lane][j]; c[j] = 0u; }
        // ---- B: count (full blocks of kMsU frames unchecked, then the tail) ----
        const size_t wstep = (size_t)kMsWarps * a.n2; // elements between two frames of this warp
        {
            const uint64_t pol_keep = ms_policy_keep();
            const float *p = col + (size_t)g * a.n2;
            int k = g;
            for (; k + (kMsU - 1) * kMsWarps < n; k += kMsWarps * kMsU) {
                unsigned raw[kMsU];
#pragma unroll
                for (int u = 0; u < kMsU; ++u) { raw[u] = ms_ld_hint(p, pol_keep); p += wstep; }
#pragma unroll
                for (int u = 0; u < kMsU; ++u) {
                    const unsigned key = ms_key(raw[u]);
#pragma unroll
                    for (int j = 0; j < NP; ++j) ms_count(c[j], key, pv[j]);
                }
            }
            for (; k < n; k += kMsWarps, p += wstep) {
                const unsigned key = ms_key(ms_ld(p));
#pragma unroll
                for (int j = 0; j < NP; ++j) ms_count(c[j], key, pv[j]);
            }
        }
        unsigned *s_c = s_list; // [kMsWarps][kMsPix][NP + 1] (the samples are no longer needed: every warp passed the barrier above)
        __syncthreads();
#pragma unroll
        for (int j = 0; j < NP; ++j) s_c[(g * kMsPix + lane) * (NP + 1) + j] = c[j];
        __syncthreads();
        if (g == 0) {
#pragma unroll
            for (int j = 0; j < NP; ++j) {
                unsigned t = 0u;
                for (int w = 0; w < kMsWarps; ++w) t += s_c[(w * kMsPix + lane) * (NP + 1) + j];
                c[j] = t;
            }
            const MsBin b = ms_pick_bin<NP>(c, pv, r0);
            s_lo[lane] = b.lo; s_hi[lane] = b.hi; s_below[lane] = b.below; s_haslo[lane] = b.has_lo ? 1u : 0u;
        }
        __syncthreads();
        lo = s_lo[lane]; hi = s_hi[lane];
    } else {
        __syncthreads();
    }
    // ---- C: collect the bin ----
    unsigned n_eq = 0u, hi_min = kFull;
    unsigned *row = s_list + lane * kMsRow;
    const unsigned cnt_addr = (unsigned)__cvta_generic_to_shared(&s_cnt[lane]), row_addr = (unsigned)__cvta_generic_to_shared(row);
    auto collect = [&](unsigned rawbits) {
        const unsigned key = ms_key(rawbits);
        const float xf = __uint_as_float(rawbits);
        nan |= xf != xf;                           // one FSETP on the raw value
        if (key == lo) ++n_eq;                     // without a lower pivot lo is 0, the key of a NaN: never counted for clean data
        if (key >= hi) hi_min = min(hi_min, key);
        // append, predicated rather than branched (the compiler's version spends a BSSY / BRA / BSYNC triple per key, and in
        // almost every warp some lane has a key in its bin): slot = atomic counter of the spaxel, stored when below the cap
        const unsigned inbin = (key > lo && key < hi) ? 1u : 0u;
        asm volatile(
            "{\n\t"
            ".reg .pred p, q;\n\t"
            ".reg .u32 at, ad;\n\t"
            "setp.ne.u32 p, %0, 0;\n\t"
            "@p atom.shared.add.u32 at, [%1], 1;\n\t"
            "setp.lt.and.u32 q, at, %2, p;\n\t"
            "mad.lo.u32 ad, at, 4, %3;\n\t"
            "@q st.shared.u32 [ad], %4;\n\t"
            "}"
            :
            : "r"(inbin), "r"(cnt_addr), "r"((unsigned)kMsCap), "r"(row_addr), "r"(key)
            : "memory");
    };
    if (active && n <= kMsCap) {
        // the whole window fits the list: frame k goes to slot k, nothing to prune, no atomics
        const size_t wstep = (size_t)kMsWarps * a.n2;
        const float *p = col + (size_t)g * a.n2;
        int k = g;
        for (; k + (kMsU - 1) * kMsWarps < n; k += kMsWarps * kMsU) {
            unsigned raw[kMsU];
#pragma unroll
            for (int u = 0; u < kMsU; ++u) { raw[u] = ms_ld(p); p += wstep; }
#pragma unroll
            for (int u = 0; u < kMsU; ++u) {
                const unsigned key = ms_key(raw[u]);
                nan |= (key - 0x00800000u) > 0xFF000000u;
                row[k + u * kMsWarps] = key;
            }
        }
        for (; k < n; k += kMsWarps, p += wstep) {
            const unsigned key = ms_key(ms_ld(p));
            nan |= (key - 0x00800000u) > 0xFF000000u;
            row[k] = key;
        }
        if (g == 0) s_cnt[lane] = (unsigned)n;
    } else if (active) {
        // Backwards: pass B swept the window upwards, so its last frames are the ones still in L2 (the resident tiles of
        // all CTAs, 296 x 32 spaxels x n frames, are about the size of L2 for n = 3840).  The list is unordered anyway.
        const size_t wstep = (size_t)kMsWarps * a.n2;
        int i = g < n ? (n - g + kMsWarps - 1) / kMsWarps : 0; // this warp's frames are k = g + j * kMsWarps, j < i
        const float *p = col + ((size_t)g + (size_t)(i > 0 ? i - 1 : 0) * kMsWarps) * a.n2;
        const uint64_t pol_last = ms_policy_last();
        for (; i > kMsUC; i -= kMsUC) { // `>`: the pointer never steps below the warp's first frame
            unsigned raw[kMsUC];
#pragma unroll
            for (int u = 0; u < kMsUC; ++u) { raw[u] = ms_ld_hint(p, pol_last); p -= wstep; }
#pragma unroll
            for (int u = 0; u < kMsUC; ++u) collect(raw[u]);
        }
        for (; i > 0; --i) {
            collect(ms_ld(p));
            if (i > 1) p -= wstep;
        }
    }
    s_eq[g][lane] = n_eq;
    s_himin[g][lane] = hi_min;
    __syncthreads();
    // ---- D: one warp per spaxel ----
    for (int p = g; p < kMsPix; p += kMsWarps) {
        const int64_t px = px0 + p;
        if (px >= a.n2) continue; // warp-uniform
        unsigned *list = s_list + p * kMsRow;
        unsigned m = s_cnt[p];
        unsigned blo = s_lo[p], bhi = s_hi[p];
        bool bhas_lo = s_haslo[p] != 0u;
        unsigned eq = lane < kMsWarps ? s_eq[lane][p] : 0u;
        eq = __reduce_add_sync(kFull, eq);
        unsigned above = lane < kMsWarps ? s_himin[lane][p] : kFull;
        above = __reduce_min_sync(kFull, above);
        unsigned r = r0 - s_below[p];
        unsigned klo = 0u, khi = 0u;
        bool done = false;
        const float *pcol = a.cube + a.start * a.n2 + px;
        unsigned seed = (unsigned)px * 0x9E3779B9u + 0x7F4A7C15u;
        if (bhas_lo && r < eq) { // the wanted key is lo itself
            klo = blo;
            khi = blo;
            if (even && !(r + 1 < eq)) {
                unsigned mn = kFull;
                if (m <= (unsigned)kMsCap) {
                    for (int i = lane; i < (int)m; i += 32) mn = min(mn, list[i]);
                } else {
                    for (int k = lane; k < n; k += 32) {
                        const unsigned key = ms_key(ms_ld(pcol + (size_t)k * a.n2));
                        if (key > blo && key < bhi) mn = min(mn, key);
                    }
                }
                mn = __reduce_min_sync(kFull, mn);
                khi = mn != kFull ? mn : above;
            }
            done = true;
        } else {
            if (bhas_lo) r -= eq;
            while (!done && m > (unsigned)kMsCap) {
                done = ms_refine(pcol, a.n2, n, list, blo, bhi, bhas_lo, r, m, above, even, lane, seed, klo, khi);
                seed = seed * 1664525u + 1013904223u;
            }
        }
        if (!done) {
            MedState st{(int)m, r, above, above != kFull};
            while (!done && st.m > 32) {
                done = med_round(list, st, even, lane, seed, klo, khi);
                seed = seed * 1664525u + 1013904223u;
            }
            if (!done) med_rank32(list, st, even, lane, klo, khi);
        }
        if (lane == 0) a.out[px] = med_finish(klo, khi, even);
    }
    if (nan && n > 1) atomicMax(a.flag, 1u); // a single frame is never compared by the sort
}

// ---------------------------------------------------------------------------------------------------------------
// median_stream1_kernel (round 2): ONE pass over the window for 768 < n <= 4400 frames (the ALMA window).
//
// median_stream_kernel reads the window twice (count, then collect): 8 B per voxel of DRAM traffic and an issue-bound count
// pass (profiles/r1_median_ncu_full.json).  Here the bracket that will hold the median is fixed BEFORE the stream:
//   A' 256 stratified sample frames per spaxel (the same frames for the whole CTA: coalesced, 6.7 % of the window), sorted
//      by a warp as 8 keys per lane (bitonic, 36 compare-exchange steps); lo / hi = the samples of rank 128 -+ 17.  The
//      sample rank of the true median has sigma ~ 7.7, so the bracket holds it for ~97 % of the spaxels and contains
//      ~13 % of the window (510 +- 80 keys: overflows the 768-slot list for < 0.1 % of the spaxels);
//   BC one stream: below += key < lo, count key == lo, track the smallest key >= hi, append lo < key < hi to the spaxel's
//      list in shared memory -- the collect pass of the two-pass kernel plus one compare;
//   D  a spaxel whose wanted rank fell outside its bracket (r0 < below or r0 >= below + eq + m) has its outer bin
//      re-collected from the global column by the WHOLE CTA (all frames in flight at once); then one warp per spaxel finishes
//      on the list exactly as before, refinement rounds for an overflowing list included: exactness never depends on the
//      data, only the time of ~3 % of the spaxels.
// 4.3 B per voxel of DRAM traffic instead of 8, and no count pass.
// ---------------------------------------------------------------------------------------------------------------
constexpr int kMs1HalfWidth = 17;          // bracket = sample ranks 128 -+ 17 of 256
constexpr int kMs1WingWidth = 40;          // a missed bracket falls back to the wing up to rank 128 -+ 40 (5 sigma)
constexpr int64_t kMs1MaxFrames = 4400;    // beyond, 15 % of the window no longer fits the list with margin

// ascending bitonic sort of 256 keys held 8 per lane: index = lane * 8 + e
__device__ __forceinline__ void ms_sort256(unsigned (&v)[8], int lane) {
#pragma unroll
    for (int k = 2; k <= 256; k <<= 1) {
#pragma unroll
        for (int j = k >> 1; j > 0; j >>= 1) {
            if (j >= 8) {
                const int lj = j >> 3;
                const bool asc = ((lane << 3) & k) == 0;
                const bool low = (lane & lj) == 0;
                const bool take_min = asc == low;
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    const unsigned o = __shfl_xor_sync(kFull, v[e], lj);
                    v[e] = take_min ? min(v[e], o) : max(v[e], o);
                }
            } else {
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    if ((e & j) == 0) {
                        const bool asc = k >= 8 ? (((lane << 3) & k) == 0) : ((e & k) == 0);
                        const unsigned x = v[e], y = v[e | j];
                        const unsigned mn = min(x, y), mx = max(x, y);
                        v[e] = asc ? mn : mx;
                        v[e | j] = asc ? mx : mn;
                    }
                }
            }
        }
    }
}

__global__ void __launch_bounds__(kMsPix * kMsWarps, 2) median_stream1_kernel(const MedianK a) {
    extern __shared__ unsigned s_list[]; // [kMsPix][kMsRow] candidate lists; phase A' aliases the samples onto it
    __shared__ unsigned s_lo[kMsPix], s_hi[kMsPix], s_lo2[kMsPix], s_hi2[kMsPix], s_cnt[kMsPix];
    __shared__ unsigned s_eq[kMsWarps][kMsPix], s_himin[kMsWarps][kMsPix], s_bel[kMsWarps][kMsPix];
    const int n = (int)a.n;
    const int tid = threadIdx.x, g = tid >> 5, lane = tid & 31;
    const int64_t px0 = (int64_t)blockIdx.x * kMsPix;
    const int64_t mypx = px0 + lane;
    const bool active = mypx < a.n2;
    const float *col = a.cube + a.start * a.n2 + (active ? mypx : a.n2 - 1); // inactive lanes read a valid address
    const unsigned r0 = (unsigned)((n - 1) / 2);
    const bool even = (n & 1) == 0;
    bool nan = false;
    long long t_tr = a.trace ? clock64() : 0;
    if (tid < kMsPix) s_cnt[tid] = 0u;
    // ---- A': 256 stratified sample frames -> bracket ----
    {
        unsigned *s_smp = s_list; // [kMsPix][257]
        const int stride = n >> 8; // n > kMsCap >= 256
        unsigned raw[256 / kMsWarps]; // this warp's 16 sample frames: all loads in flight before the first store
#pragma unroll
        for (int u = 0; u < 256 / kMsWarps; ++u) {
            const int i = g + u * kMsWarps;
            unsigned h = ((unsigned)blockIdx.x * 0x9E3779B9u + (unsigned)i * 0x85EBCA6Bu) * 0xC2B2AE35u;
            h ^= h >> 16;
            const int f = (int)(((long long)i * n) >> 8) + (int)(h % (unsigned)stride);
            raw[u] = ms_ld(col + (size_t)f * a.n2);
        }
        // the low 8 bits are replaced by the sample's index: the 256 values are distinct, so lo < hi strictly even for a
        // constant spectrum (a pivot does not have to be a key of the spectrum)
#pragma unroll
        for (int u = 0; u < 256 / kMsWarps; ++u) s_smp[lane * 257 + g + u * kMsWarps] = (ms_key(raw[u]) & ~255u) | (unsigned)(g + u * kMsWarps);
        __syncthreads();
        ms_trace(a, 0, t_tr);
        for (int p = g; p < kMsPix; p += kMsWarps) {
            unsigned v[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) v[e] = s_smp[p * 257 + lane * 8 + e];
            ms_sort256(v, lane);
            constexpr int RL = 128 - kMs1HalfWidth, RH = 128 + kMs1HalfWidth, RL2 = 128 - kMs1WingWidth, RH2 = 128 + kMs1WingWidth;
            const unsigned lo_s = __shfl_sync(kFull, v[RL & 7], RL >> 3), hi_s = __shfl_sync(kFull, v[RH & 7], RH >> 3);
            const unsigned lo2_s = __shfl_sync(kFull, v[RL2 & 7], RL2 >> 3), hi2_s = __shfl_sync(kFull, v[RH2 & 7], RH2 >> 3);
            if (lane == 0) { s_lo[p] = lo_s; s_hi[p] = hi_s; s_lo2[p] = lo2_s; s_hi2[p] = hi2_s; }
        }
        __syncthreads(); // the samples are dead: the lists may be written
        ms_trace(a, 1, t_tr);
    }
    const unsigned lo = s_lo[lane], hi = s_hi[lane];
    // ---- BC: one stream over the window ----
    unsigned n_eq = 0u, hi_min = kFull, below = 0u;
    unsigned *row = s_list + lane * kMsRow;
    const unsigned cnt_addr = (unsigned)__cvta_generic_to_shared(&s_cnt[lane]), row_addr = (unsigned)__cvta_generic_to_shared(row);
    auto collect = [&](unsigned rawbits) {
        const unsigned key = ms_key(rawbits);
        const float xf = __uint_as_float(rawbits);
        nan |= xf != xf;
        ms_count(below, key, lo);
        if (key == lo) ++n_eq;
        if (key >= hi) hi_min = min(hi_min, key);
        const unsigned inbin = (key > lo && key < hi) ? 1u : 0u;
        asm volatile(
            "{\n\t"
            ".reg .pred p, q;\n\t"
            ".reg .u32 at, ad;\n\t"
            "setp.ne.u32 p, %0, 0;\n\t"
            "@p atom.shared.add.u32 at, [%1], 1;\n\t"
            "setp.lt.and.u32 q, at, %2, p;\n\t"
            "mad.lo.u32 ad, at, 4, %3;\n\t"
            "@q st.shared.u32 [ad], %4;\n\t"
            "}"
            :
            : "r"(inbin), "r"(cnt_addr), "r"((unsigned)kMsCap), "r"(row_addr), "r"(key)
            : "memory");
    };
    if (active) {
        const size_t wstep = (size_t)kMsWarps * a.n2;
        const float *p = col + (size_t)g * a.n2;
        int k = g;
        for (; k + (kMsUC - 1) * kMsWarps < n; k += kMsWarps * kMsUC) {
            unsigned raw[kMsUC];
#pragma unroll
            for (int u = 0; u < kMsUC; ++u) { raw[u] = ms_ld(p); p += wstep; } // default L2 policy: a re-collect (D1) finds most of the tile still cached
#pragma unroll
            for (int u = 0; u < kMsUC; ++u) collect(raw[u]);
        }
        for (; k < n; k += kMsWarps, p += wstep) collect(ms_ld(p));
    }
    s_eq[g][lane] = n_eq;
    s_himin[g][lane] = hi_min;
    s_bel[g][lane] = below;
    __syncthreads();
    ms_trace(a, 2, t_tr);
    // ---- D0: lane = spaxel: does the bracket hold the wanted rank?  If not, the bin becomes the WING next to it, bounded by
    //      the samples of rank 128 -+ 40 (5 sigma): ~9 % of the window, it fits the list ----
    __shared__ unsigned s_r[kMsPix], s_eqt[kMsPix], s_abv[kMsPix], s_haslo[kMsPix], s_miss[kMsPix], s_belt[kMsPix];
    if (g == 0) {
        unsigned eq = 0u, bel = 0u, above = kFull;
#pragma unroll
        for (int w = 0; w < kMsWarps; ++w) { eq += s_eq[w][lane]; bel += s_bel[w][lane]; above = min(above, s_himin[w][lane]); }
        const unsigned m = s_cnt[lane];
        unsigned r, miss = 0u;
        if (r0 < bel) { // below the bracket: (lo2, lo); the rank inside it is known once the wing has been counted
            s_hi[lane] = lo;
            s_lo[lane] = s_lo2[lane];
            r = r0;
            miss = 1u;
        } else if (r0 - bel >= eq + m) { // above it: [hi, hi2), hi itself counted as the new `eq`
            r = r0 - bel - eq - m;
            s_lo[lane] = hi;
            s_hi[lane] = s_hi2[lane];
            miss = 2u;
        } else {
            r = r0 - bel;
        }
        if (miss) { s_cnt[lane] = 0u; eq = 0u; above = kFull; }
        s_r[lane] = r; s_eqt[lane] = eq; s_abv[lane] = above; s_haslo[lane] = 1u; s_belt[lane] = bel;
        s_miss[lane] = active ? miss : 0u;
        if (a.trace) {
            if (active && miss) atomicAdd(a.trace + 5, 1ull);
            if (active && !miss && m > (unsigned)kMsCap) atomicAdd(a.trace + 7, 1ull);
            if (active) atomicAdd(a.trace + 6, (unsigned long long)m);
        }
    }
    __syncthreads();
    // ---- D1: the whole CTA re-collects the bin of a spaxel whose bracket missed (every frame of the column in flight at
    //      once: a few microseconds, where one warp walking the column would stall the CTA for ~100) ----
    auto recollect = [&]() {
        for (int p = 0; p < kMsPix; ++p) {
            if (!s_miss[p]) continue; // CTA-uniform
            const float *pcol = a.cube + a.start * a.n2 + (px0 + p);
            const unsigned plo = s_lo[p], phi = s_hi[p];
            const bool phas = s_haslo[p] != 0u;
            unsigned *list = s_list + p * kMsRow;
            unsigned eqc = 0u, hm = kFull;
            constexpr int kPer = (int)((kMs1MaxFrames + kMsPix * kMsWarps - 1) / (kMsPix * kMsWarps)); // frames per thread
            unsigned kk[kPer];
#pragma unroll
            for (int u = 0; u < kPer; ++u) {
                const int k = tid + u * kMsPix * kMsWarps;
                kk[u] = k < n ? ms_key(ms_ld(pcol + (size_t)k * a.n2)) : plo; // plo: counted nowhere below
            }
#pragma unroll
            for (int u = 0; u < kPer; ++u) {
                const bool valid = tid + u * kMsPix * kMsWarps < n;
                const unsigned key = kk[u];
                if (valid && phas && key == plo) ++eqc;
                if (valid && key >= phi) hm = min(hm, key);
                // one shared atomic per warp and step, not per key: every append of this loop goes to the same counter
                const bool keep = valid && key > plo && key < phi;
                const unsigned bal = __ballot_sync(kFull, keep);
                if (bal) {
                    unsigned base = 0u;
                    if (lane == __ffs(bal) - 1) base = atomicAdd(&s_cnt[p], (unsigned)__popc(bal));
                    base = __shfl_sync(kFull, base, __ffs(bal) - 1);
                    const unsigned at = base + __popc(bal & ((1u << lane) - 1u));
                    if (keep && at < (unsigned)kMsCap) list[at] = key;
                }
            }
            eqc = __reduce_add_sync(kFull, eqc);
            hm = __reduce_min_sync(kFull, hm);
            if (lane == 0) {
                if (eqc) atomicAdd(&s_eqt[p], eqc);
                if (hm != kFull) atomicMin(&s_abv[p], hm);
            }
        }
    };
    recollect();
    __syncthreads();
    // the wing is counted: rank inside it, or (5 sigma: ~1e-7 of the spaxels) the open bin beyond it
    if (g == 0) {
        const unsigned miss = s_miss[lane];
        unsigned again = 0u;
        if (miss == 1u) {
            const unsigned below2 = s_belt[lane] - s_eqt[lane] - s_cnt[lane]; // keys < lo2
            if (r0 >= below2) {
                s_r[lane] = r0 - below2;
            } else {
                s_hi[lane] = s_lo[lane];
                s_lo[lane] = 0u;
                s_haslo[lane] = 0u;
                s_r[lane] = r0;
                again = 1u;
            }
        } else if (miss == 2u) {
            const unsigned inside = s_eqt[lane] + s_cnt[lane]; // keys in [hi, hi2)
            if (s_r[lane] >= inside) {
                s_r[lane] -= inside;
                s_lo[lane] = s_hi[lane];
                s_hi[lane] = kFull;
                again = 1u;
            }
        }
        if (again) { s_cnt[lane] = 0u; s_eqt[lane] = 0u; s_abv[lane] = kFull; }
        s_miss[lane] = again;
    }
    __syncthreads();
    recollect();
    __syncthreads();
    ms_trace(a, 3, t_tr);
    // ---- D2: one warp per spaxel ----
    for (int p = g; p < kMsPix; p += kMsWarps) {
        const int64_t px = px0 + p;
        if (px >= a.n2) continue; // warp-uniform
        unsigned *list = s_list + p * kMsRow;
        unsigned m = s_cnt[p];
        unsigned blo = s_lo[p], bhi = s_hi[p];
        bool bhas_lo = s_haslo[p] != 0u;
        const unsigned eq = s_eqt[p];
        unsigned above = s_abv[p];
        unsigned r = s_r[p];
        const float *pcol = a.cube + a.start * a.n2 + px;
        unsigned klo = 0u, khi = 0u;
        bool done = false;
        unsigned seed = (unsigned)px * 0x9E3779B9u + 0x7F4A7C15u;
        if (bhas_lo && r < eq) { // the wanted key is lo itself
            klo = blo;
            khi = blo;
            if (even && !(r + 1 < eq)) {
                unsigned mn = kFull;
                if (m <= (unsigned)kMsCap) {
                    for (int i = lane; i < (int)m; i += 32) mn = min(mn, list[i]);
                } else {
                    for (int k = lane; k < n; k += 32) {
                        const unsigned key = ms_key(ms_ld(pcol + (size_t)k * a.n2));
                        if (key > blo && key < bhi) mn = min(mn, key);
                    }
                }
                mn = __reduce_min_sync(kFull, mn);
                khi = mn != kFull ? mn : above;
            }
            done = true;
        } else {
            if (bhas_lo) r -= eq;
            while (!done && m > (unsigned)kMsCap) {
                done = ms_refine(pcol, a.n2, n, list, blo, bhi, bhas_lo, r, m, above, even, lane, seed, klo, khi);
                seed = seed * 1664525u + 1013904223u;
            }
        }
        if (!done) {
            MedState st{(int)m, r, above, above != kFull};
            while (!done && st.m > 32) {
                done = med_round(list, st, even, lane, seed, klo, khi);
                seed = seed * 1664525u + 1013904223u;
            }
            if (!done) med_rank32(list, st, even, lane, klo, khi);
        }
        if (lane == 0) a.out[px] = med_finish(klo, khi, even);
    }
    if (a.trace) { __syncthreads(); ms_trace(a, 4, t_tr); }
    if (nan && n > 1) atomicMax(a.flag, 1u);
}

// ---------------------------------------------------------------------------------------------------------------
// median_stream2_kernel (round 2): the one-pass kernel rebuilt around what its phase trace showed
// (profiles/r2_median_single_pass.json: stream 51 % of the time at 40 % issue utilisation, finish 19 %, re-collect 16 %,
// sample sort 10 %; profiles/r2_median_v2_phases.txt for the steps in between).
//
// A CTA is 8 warps (2 CTAs = 512 threads per SM, up to 128 registers each).  Thread (warp g, lane) owns the frames
// g, g+8, g+16, ... of spaxel `lane` ("its" frames, t = 0, 1, ...) and a PRIVATE segment of 100 slots of the spaxel's
// 800-slot row, in every phase:
//   S  The samples are part of the stream, not an extra read: the thread's frames t = 0, ss, 2 ss, ... (ss = ceil(frames /
//      64): ~60 per thread, ~480 per spaxel, stratified over the window) are loaded first and stay in registers.  Each
//      thread sorts ITS samples with a bitonic network in registers (672 compare-exchanges of two FMNMX, no shuffles, no
//      shared memory: 1/8 of the instructions of the warp-wide sort of 256 keys per spaxel it replaces, with twice the
//      samples) and reads the quantiles 0.5 -+ w and 0.5 -+ w2 off its run (interpolated); lo / hi / the wing pivots of the
//      spaxel are the means of the 8 runs' quantiles.  A pivot does not have to be a value of the spectrum; w is 3 sigma
//      of that estimator (host side), so the bracket misses ~0.3 % of the spaxels instead of 1.8 % and the list is no longer
//      than before (12 % of the window).
//   C  No atomics in the stream: an append is one predicated shared store into the private segment and a pointer bump
//      (the shared-atomic round trip per key, whose returned slot serialised every step, is gone).  Float compares on the
//      raw values (x == lo, lo < x < hi, x >= hi; x < lo is what remains): no key transform per voxel, NaNs are caught by
//      one min.NaN per voxel: 11 instructions per voxel against ~19.  +-0 compare equal, which only matters for the sign
//      of a zero result -- median_zero_fix_kernel decides that for every kernel of this file.  The sorted samples are
//      classified from their registers; the other frames are walked column by column (t = j ss + c for c = 1 .. ss-1) in
//      batches of 16 loads, two batches in ping-pong.
//   D  As before: a bracket that missed the wanted rank is re-collected by the whole CTA from the wing next to it (or the
//      open bin beyond when the wing pivot is not distinct); so is a spaxel with an overflowing segment.  A list that
//      still overflows is refined against the global column.  Exactness never depends on the data.
// ---------------------------------------------------------------------------------------------------------------
constexpr int kM2Warps = 8;
constexpr int kM2Seg = kMsCap / kM2Warps; // 100 private slots per (warp, spaxel)
constexpr int kM2Smp = 64;                // samples per thread (at most)
static_assert(kM2Seg * kM2Warps == kMsCap, "segments tile the row");
static_assert((size_t)kM2Warps * kM2Smp * kMsPix <= (size_t)kMsPix * kMsRow, "the sorted runs alias the list area");

__device__ __forceinline__ float m2_min_nan(float a, float b) {
    float r;
    asm("min.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
    return r;
}

// ascending bitonic sort of N values in registers (every index is a compile-time constant)
template <int N> __device__ __forceinline__ void m2_sort_regs(float (&v)[N]) {
#pragma unroll
    for (int k = 2; k <= N; k <<= 1) {
#pragma unroll
        for (int j = k >> 1; j > 0; j >>= 1) {
#pragma unroll
            for (int i = 0; i < N; ++i) {
                const int l = i ^ j;
                if (l > i) {
                    const bool asc = (i & k) == 0;
                    const float x = v[i], y = v[l];
                    const float mn = fminf(x, y), mx = fmaxf(x, y);
                    v[i] = asc ? mn : mx;
                    v[l] = asc ? mx : mn;
                }
            }
        }
    }
}

// quantile q of a sorted run of ms values in shared memory (stride 32 words): linear interpolation between two order
// statistics at the plotting position q ms - 1/2; infinities never produce a NaN
__device__ __forceinline__ float m2_run_quantile(const float *run, int ms, float q) {
    float pos = __fsub_rn(__fmul_rn(q, (float)ms), 0.5f);
    pos = fminf(fmaxf(pos, 0.0f), (float)(ms - 1));
    int i0 = (int)pos;
    if (i0 > ms - 2) i0 = ms - 2;
    if (i0 < 0) i0 = 0;
    const float fr = __fsub_rn(pos, (float)i0);
    const float x = run[i0 * 32], y = run[(ms > 1 ? i0 + 1 : i0) * 32];
    float v = __fadd_rn(__fmul_rn(__fsub_rn(1.0f, fr), x), __fmul_rn(fr, y));
    if (v != v) v = fr < 0.5f ? x : y;
    return v;
}

template <int SS> // sample stride: a thread's frames t = 0, SS, 2 SS, ... are its samples; compile-time so that a batch is whole blocks of SS - 1 frames
__global__ void __launch_bounds__(kMsPix * kM2Warps, 2) median_stream2_kernel(const MedianK a, float wq, float wq2) {
    extern __shared__ unsigned s_list[]; // [kMsPix][kMsRow]: 8 private segments of 100 slots per spaxel; phase S aliases the sorted runs onto it
    __shared__ float s_q[4][kM2Warps][kMsPix]; // per-run quantiles: lo, hi, wing lo, wing hi
    __shared__ float s_lo[kMsPix], s_hi[kMsPix];
    __shared__ unsigned s_cnt[kMsPix];
    __shared__ unsigned s_seg[kM2Warps][kMsPix], s_eq[kM2Warps][kMsPix], s_bel[kM2Warps][kMsPix];
    __shared__ unsigned s_r[kMsPix], s_eqt[kMsPix], s_abv[kMsPix], s_haslo[kMsPix], s_hashi[kMsPix], s_miss[kMsPix], s_belt[kMsPix], s_packed[kMsPix];
    const int n = (int)a.n;
    const int tid = threadIdx.x, g = tid >> 5, lane = tid & 31;
    const int64_t px0 = (int64_t)blockIdx.x * kMsPix;
    const int64_t mypx = px0 + lane;
    const bool active = mypx < a.n2;
    const float *col = a.cube + a.start * a.n2 + (active ? mypx : a.n2 - 1); // inactive lanes do the same work on a valid address
    const unsigned r0 = (unsigned)((n - 1) / 2);
    const bool even = (n & 1) == 0;
    const float kInf = __int_as_float(0x7f800000);
    if (px0 >= a.n2) return; // the grid is padded to whole clusters
    long long t_tr = a.trace ? clock64() : 0;
    // this thread's frames: t = 0 .. mine-1 is frame g + 8 t, in blocks of SS: the first frame of block j (t = j SS) is a sample
    // (j < ms <= 64: the host picks SS), its other SS - 1 frames are streamed
    const int mine = (n - g + kM2Warps - 1) / kM2Warps;
    const int ms = (mine + SS - 1) / SS;
    const size_t tstep = (size_t)kM2Warps * a.n2;   // elements between two of this thread's frames
    const float *const tcol = col + (size_t)g * a.n2;
    float nan_acc = 0.0f;
    // ---- S: samples -> sorted run -> quantiles -> bracket ----
    float smp[kM2Smp];
    {
        const float *sp = tcol;
#pragma unroll
        for (int j0 = 0; j0 < kM2Smp; j0 += 16) {
            if (j0 + 16 <= ms) {
#pragma unroll
                for (int j = j0; j < j0 + 16; ++j) { smp[j] = __uint_as_float(ms_ld(sp)); sp += (size_t)SS * tstep; }
            } else {
#pragma unroll
                for (int j = j0; j < j0 + 16; ++j) {
                    smp[j] = kInf; // pads sort behind the samples: the first ms sorted values are the samples
                    if (j < ms) smp[j] = __uint_as_float(ms_ld(sp));
                    sp += (size_t)SS * tstep;
                }
            }
        }
    }
#pragma unroll
    for (int j = 0; j < kM2Smp; ++j) nan_acc = m2_min_nan(nan_acc, smp[j]);
    ms_trace(a, 0, t_tr);
    m2_sort_regs<kM2Smp>(smp);
    {
        float *run = reinterpret_cast<float *>(s_list) + (size_t)g * kM2Smp * 32 + lane; // [g][j][lane]: conflict-free
#pragma unroll
        for (int j = 0; j < kM2Smp; ++j) run[j * 32] = smp[j];
        s_q[0][g][lane] = m2_run_quantile(run, ms, 0.5f - wq);
        s_q[1][g][lane] = m2_run_quantile(run, ms, 0.5f + wq);
        s_q[2][g][lane] = m2_run_quantile(run, ms, 0.5f - wq2);
        s_q[3][g][lane] = m2_run_quantile(run, ms, 0.5f + wq2);
    }
    __syncthreads(); // the runs are dead: the lists may be written
    float lo = 0.0f, hi = 0.0f;
#pragma unroll
    for (int w = 0; w < kM2Warps; ++w) {
        lo = __fadd_rn(lo, __fmul_rn(s_q[0][w][lane], 0.125f));
        hi = __fadd_rn(hi, __fmul_rn(s_q[1][w][lane], 0.125f));
    }
    if (lo != lo) lo = s_q[0][0][lane]; // +inf and -inf among the runs
    if (hi != hi) hi = s_q[1][0][lane];
    // the classes below must be disjoint: lo < hi strictly.  Tied quantiles: the bracket is everything above lo (the wanted
    // rank is then almost surely in the `== lo` class); nothing is above +inf.
    if (!(hi > lo)) hi = (lo == kInf) ? __int_as_float(0x7fc00000) : kInf;
    ms_trace(a, 1, t_tr);
    // ---- C: classify.  x == lo, lo < x < hi (appended), x >= hi (counted); x < lo is what remains.
    //      The stream is bound by the ALU pipe (compares, min/max and integer adds issue at half rate: profiles/
    //      r2_median_v2_phases.txt), so a voxel costs 5 ALU instructions on the fast path -- three compares (the third takes
    //      !(x >= hi) as a predicate operand), the list pointer and the load address -- and the two counters are predicated FADDs
    //      on the FMA pipe (exact: at most 550 frames per thread).  Two things the stream no longer does per voxel:
    //        * NaN detection: the third compare is "greater or unordered", so a NaN is APPENDED to the segment (the fast path
    //          has room for a whole batch by construction) and found when the thread scans its <= 100 slots after the stream;
    //          the checked path (segment nearly full, batch tails) keeps its min.NaN.  A CTA that saw a NaN sets the flag and
    //          stops: a candidate list never holds a NaN.
    //        * the smallest value >= hi (the even-count successor when the wanted rank is the bracket's last key): that is one
    //          spaxel in thousands, known in D0 from the counts, and served by the re-collect. ----
    float f_eq_cnt = 0.0f, f_ge_cnt = 0.0f;
    const unsigned seg0 = (unsigned)__cvta_generic_to_shared(s_list + lane * kMsRow + g * kM2Seg);
    unsigned wp = seg0; // keeps counting past the segment: D0 sees the overflow
    const unsigned wend = seg0 + kM2Seg * 4u;
    auto classify_checked = [&](unsigned rawbits) {
        asm volatile(
            "{\n\t"
            ".reg .pred pe, pg, pi, ps;\n\t"
            ".reg .f32 x;\n\t"
            "mov.b32 x, %4;\n\t"
            "setp.eq.f32 pe, x, %5;\n\t"
            "@pe add.f32 %0, %0, 0f3F800000;\n\t"
            "setp.ge.f32 pg, x, %6;\n\t"
            "@pg add.f32 %1, %1, 0f3F800000;\n\t"
            "setp.gt.and.f32 pi, x, %5, !pg;\n\t"
            "setp.lt.and.u32 ps, %3, %7, pi;\n\t"
            "@ps st.shared.u32 [%3], %4;\n\t"
            "@pi add.u32 %3, %3, 4;\n\t"
            "min.NaN.f32 %2, %2, x;\n\t"
            "}"
            : "+f"(f_eq_cnt), "+f"(f_ge_cnt), "+f"(nan_acc), "+r"(wp)
            : "r"(rawbits), "f"(lo), "f"(hi), "r"(wend)
            : "memory");
    };
    auto classify = [&](unsigned rawbits) { // the caller has checked that 16 more appends fit the segment
        asm volatile(
            "{\n\t"
            ".reg .pred pe, pg, pi;\n\t"
            ".reg .f32 x;\n\t"
            "mov.b32 x, %3;\n\t"
            "setp.eq.f32 pe, x, %4;\n\t"
            "@pe add.f32 %0, %0, 0f3F800000;\n\t"
            "setp.ge.f32 pg, x, %5;\n\t"
            "@pg add.f32 %1, %1, 0f3F800000;\n\t"
            "setp.gtu.and.f32 pi, x, %4, !pg;\n\t" // greater or unordered: a NaN goes to the list
            "@pi st.shared.u32 [%2], %3;\n\t"
            "@pi add.u32 %2, %2, 4;\n\t"
            "}"
            : "+f"(f_eq_cnt), "+f"(f_ge_cnt), "+r"(wp)
            : "r"(rawbits), "f"(lo), "f"(hi)
            : "memory");
    };
#pragma unroll
    for (int j0 = 0; j0 < kM2Smp; j0 += 16) { // the sorted samples, from their registers
        if (j0 + 16 <= ms && wp + 4u * 16u <= wend) {
#pragma unroll
            for (int j = j0; j < j0 + 16; ++j) classify(__float_as_uint(smp[j]));
        } else {
#pragma unroll
            for (int j = j0; j < j0 + 16; ++j)
                if (j < ms) classify_checked(__float_as_uint(smp[j]));
        }
    }
    // the segment starts with this thread's samples of the bracket, in ascending order (they were classified sorted): the first
    // round of the finish picks its pivots there
    const unsigned ks = min((wp - seg0) >> 2, (unsigned)kM2Seg);
    {
        // the other frames: a batch is kNB whole blocks (kBatch = kNB (SS - 1) ~ 16 loads, every address one 64-bit add from the
        // last), two batches in ping-pong; the last blocks, which may be cut by the end of the window, are one guarded batch
        constexpr int kMem = SS - 1;
        constexpr int kNB = (16 + kMem / 2) / kMem;
        constexpr int kBatch = kNB * kMem;
        const int nfb = mine / SS;     // blocks with all their frames inside the window
        int left = nfb / kNB;          // full batches
        const float *p = tcol + tstep; // block 0, frame 1
        auto load = [&](unsigned (&raw)[kBatch]) {
#pragma unroll
            for (int b = 0; b < kNB; ++b) {
#pragma unroll
                for (int c = 0; c < kMem; ++c) {
                    raw[b * kMem + c] = ms_ld(p);
                    p += c == kMem - 1 ? 2 * tstep : tstep; // over the next block's sample
                }
            }
        };
        auto process = [&](const unsigned (&raw)[kBatch]) {
            if (wp + 4u * kBatch <= wend) {
#pragma unroll
                for (int u = 0; u < kBatch; ++u) classify(raw[u]);
            } else {
#pragma unroll
                for (int u = 0; u < kBatch; ++u) classify_checked(raw[u]);
            }
        };
        const int t_tail = (left * kNB) * SS; // first frame of the first block that is not part of a full batch
        unsigned ra[kBatch], rb[kBatch];
        if (left > 0) load(ra);
        while (left > 0) {
            if (left > 1) load(rb);
            process(ra);
            if (--left == 0) break;
            if (left > 1) load(ra);
            process(rb);
            --left;
        }
        // tail: at most kNB blocks (the last one possibly short), all loads in flight at once
        unsigned rt[kBatch + kMem];
        const float *q = tcol + (size_t)t_tail * tstep;
#pragma unroll
        for (int b = 0; b < kNB + 1; ++b) {
#pragma unroll
            for (int c = 0; c < kMem; ++c) {
                const int t = t_tail + b * SS + 1 + c;
                rt[b * kMem + c] = 0u;
                if (t < mine) rt[b * kMem + c] = ms_ld(q + (size_t)(b * SS + 1 + c) * tstep);
            }
        }
#pragma unroll
        for (int b = 0; b < kNB + 1; ++b) {
#pragma unroll
            for (int c = 0; c < kMem; ++c)
                if (t_tail + b * SS + 1 + c < mine) classify_checked(rt[b * kMem + c]);
        }
    }
    const unsigned cnt = (wp - seg0) >> 2;
    {
        // NaNs the fast path appended: this thread's own stores, no barrier needed
        const float *own = reinterpret_cast<const float *>(s_list + lane * kMsRow + g * kM2Seg);
        const unsigned c = min(cnt, (unsigned)kM2Seg);
        for (unsigned i = 0; i < c; ++i) nan_acc = m2_min_nan(nan_acc, own[i]);
    }
    const bool nan = nan_acc != nan_acc;
    const unsigned n_eq = (unsigned)f_eq_cnt, n_ge = (unsigned)f_ge_cnt;
    s_seg[g][lane] = cnt;
    s_eq[g][lane] = n_eq;
    s_bel[g][lane] = (unsigned)mine - n_eq - n_ge - cnt; // x < lo
    if (tid < kMsPix) { s_lo[tid] = 0.0f; s_hi[tid] = 0.0f; }
    // a NaN anywhere in the tile: the reference panics (the host reports it), the tile's results do not matter -- and no list may
    // hold a NaN when the finish runs (its rounds would never drop it)
    if (__syncthreads_or(nan && active)) {
        if (nan && active && n > 1) atomicMax(a.flag, 1u);
        return;
    }
    ms_trace(a, 2, t_tr);
    // ---- D0: lane = spaxel: does the bracket hold the wanted rank?  If not, the bin becomes the WING next to it (it fits the
    //      list), or the open bin beyond when the wing pivot is not distinct from the bracket's ----
    // s_miss: 0 hit, segmented list; 1 wing below; 2 wing above; 3 hit, but a segment overflowed: same bracket, re-collected
    //         into a contiguous list; 4 / 5 open bin below / above (final)
    if (g == 0) {
        unsigned eq = 0u, bel = 0u, mt = 0u;
        float lo2 = 0.0f, hi2 = 0.0f;
        bool ovf = false;
#pragma unroll
        for (int w = 0; w < kM2Warps; ++w) {
            const unsigned c = s_seg[w][lane];
            eq += s_eq[w][lane];
            bel += s_bel[w][lane];
            mt += c;
            ovf |= c > (unsigned)kM2Seg;
            lo2 = __fadd_rn(lo2, __fmul_rn(s_q[2][w][lane], 0.125f));
            hi2 = __fadd_rn(hi2, __fmul_rn(s_q[3][w][lane], 0.125f));
        }
        unsigned r, miss = 0u, haslo = 1u, hashi = 1u;
        float blo = lo, bhi = hi;
        if (r0 < bel) {
            r = r0;
            bhi = lo;
            if (lo2 < lo) { blo = lo2; miss = 1u; }
            else { haslo = 0u; miss = 4u; }
        } else if (r0 - bel >= eq + mt) { // above it: [hi, hi2), hi itself counted as the new `eq`
            r = r0 - bel - eq - mt;
            blo = hi;
            if (hi2 > hi) { bhi = hi2; miss = 2u; }
            else { hashi = 0u; miss = 5u; }
        } else {
            r = r0 - bel;
            // an overflowing segment only matters when the list is needed: not when the answer is lo itself
            const bool answer_is_lo = r < eq && (!even || r + 1 < eq);
            // even count and the wanted rank is the last value below hi: its successor is the smallest value >= hi, which the
            // stream does not track -- the re-collect of the same bracket finds it
            const bool need_above = even && r0 + 1u == bel + eq + mt;
            if ((ovf && !answer_is_lo) || need_above) miss = 3u;
            if (ovf && answer_is_lo && !need_above) mt = 0u;
            if (bhi != bhi) hashi = 0u; // lo == +inf: nothing above it
        }
        unsigned above = kFull; // becomes known when the finish drops candidates above a pivot, or through the re-collect
        if (miss) { mt = 0u; eq = 0u; }
        s_lo[lane] = blo; s_hi[lane] = bhi;
        s_cnt[lane] = mt;
        s_r[lane] = r; s_eqt[lane] = eq; s_abv[lane] = above; s_haslo[lane] = haslo; s_hashi[lane] = hashi; s_belt[lane] = bel;
        s_miss[lane] = active ? miss : 0u;
        s_packed[lane] = (active && miss) || (ovf && !miss) ? 1u : 0u; // the list is contiguous keys already (re-collected), or not needed
        if (a.trace) {
            if (active && miss && miss != 3u) atomicAdd(a.trace + 5, 1ull);
            if (active && miss == 3u) atomicAdd(a.trace + 7, 1ull);
            if (active && !miss) atomicAdd(a.trace + 6, (unsigned long long)mt);
        }
    }
    // most CTAs have nothing to re-collect: one barrier that also carries the vote
    const bool any_miss = __syncthreads_or(g == 0 && s_miss[lane] != 0u) != 0;
    // ---- D1: the whole CTA re-collects the bin of a spaxel whose bracket missed (every frame of the column in flight) ----
    auto recollect = [&]() {
        for (int p = 0; p < kMsPix; ++p) {
            if (!s_miss[p]) continue; // CTA-uniform
            const float *pcol = a.cube + a.start * a.n2 + (px0 + p);
            const float plo = s_lo[p], phi = s_hi[p];
            const bool phas = s_haslo[p] != 0u, phashi = s_hashi[p] != 0u;
            unsigned *list = s_list + p * kMsRow;
            unsigned eqc = 0u, hm = kFull;
            constexpr int kPer = (int)((kMs1MaxFrames + kMsPix * kM2Warps - 1) / (kMsPix * kM2Warps)); // frames per thread
            unsigned kk[kPer];
#pragma unroll
            for (int u = 0; u < kPer; ++u) {
                const int k = tid + u * kMsPix * kM2Warps;
                kk[u] = k < n ? ms_ld(pcol + (size_t)k * a.n2) : 0u;
            }
#pragma unroll
            for (int u = 0; u < kPer; ++u) {
                const bool valid = tid + u * kMsPix * kM2Warps < n;
                const float x = __uint_as_float(kk[u]);
                const unsigned key = ms_key(kk[u]);
                if (valid && phas && x == plo) ++eqc;
                if (valid && phashi && x >= phi) hm = min(hm, key);
                // one shared atomic per warp and step, not per key: every append of this loop goes to the same counter
                const bool keep = valid && (!phas || x > plo) && (!phashi || x < phi);
                const unsigned bal = __ballot_sync(kFull, keep);
                if (bal) {
                    unsigned base = 0u;
                    if (lane == __ffs(bal) - 1) base = atomicAdd(&s_cnt[p], (unsigned)__popc(bal));
                    base = __shfl_sync(kFull, base, __ffs(bal) - 1);
                    const unsigned at = base + __popc(bal & ((1u << lane) - 1u));
                    if (keep && at < (unsigned)kMsCap) list[at] = key;
                }
            }
            eqc = __reduce_add_sync(kFull, eqc);
            hm = __reduce_min_sync(kFull, hm);
            if (lane == 0) {
                if (eqc) atomicAdd(&s_eqt[p], eqc);
                if (hm != kFull) atomicMin(&s_abv[p], hm);
            }
        }
    };
    if (any_miss) {
        recollect();
        __syncthreads();
        // the wing is counted: rank inside it, or (6 sigma) the open bin beyond it
        if (g == 0) {
            const unsigned miss = s_miss[lane];
            unsigned again = 0u;
            if (miss == 1u) {
                const unsigned below2 = s_belt[lane] - s_eqt[lane] - s_cnt[lane]; // values < lo2
                if (r0 >= below2) {
                    s_r[lane] = r0 - below2;
                } else {
                    s_hi[lane] = s_lo[lane];
                    s_haslo[lane] = 0u;
                    s_r[lane] = r0;
                    again = 4u;
                }
            } else if (miss == 2u) {
                const unsigned inside = s_eqt[lane] + s_cnt[lane]; // values in [hi, hi2)
                if (s_r[lane] >= inside) {
                    s_r[lane] -= inside;
                    s_lo[lane] = s_hi[lane];
                    s_hashi[lane] = 0u;
                    again = 5u;
                }
            }
            if (again) { s_cnt[lane] = 0u; s_eqt[lane] = 0u; s_abv[lane] = kFull; }
            s_miss[lane] = again;
        }
        __syncthreads();
        recollect();
        __syncthreads();
    }
    ms_trace(a, 3, t_tr);
    // ---- F: the finish.  The candidates S of a spaxel (values strictly inside its bin) stay where they are, spread over the 8
    //      private segments, and all 256 threads work on all 32 spaxels at once in the (warp g, lane = spaxel) layout of the
    //      stream: a round = every thread offers one of its candidates as a pivot; every thread counts its segment against the
    //      spaxel's 8 sorted pivots (registers, no shuffles), the counts are summed with shared-memory reductions; every thread
    //      then compacts ITS segment in place to the bin that holds the wanted rank (no data moves between threads).  ~5x
    //      fewer candidates per round; when <= 32 are left they are gathered and ranked by counting.  One warp per spaxel
    //      (the earlier finish) had 2 x 8 warps per SM waiting on dependent shuffles and ballots: 56k of 165k cycles per CTA. ----
    float *const f_pv = &s_q[0][0][0];                                  // [8][32] pivot offers
    unsigned *const f_c = reinterpret_cast<unsigned *>(&s_q[1][0][0]);  // [2][32][8] pivot counts, two rounds' worth
    unsigned *const f_k = reinterpret_cast<unsigned *>(&s_q[3][0][0]);  // [8][32] candidates per thread
    unsigned *const f_mode = s_miss;  // 0 live: klo = S[r] (and khi = S[r+1] or `above`); 1 live: klo is known, khi = min S; 2 done; 3 slow path
    unsigned *const f_klo = s_haslo, *const f_khi = s_hashi; // results as keys
    unsigned *const f_eq = s_eqt, *const f_m2 = s_belt;      // per-round sums: candidates equal to the bin's lower pivot, candidates kept
    float *const seg = reinterpret_cast<float *>(s_list) + lane * kMsRow + g * kM2Seg;
    unsigned c_own;
    {
        // a re-collected list is contiguous keys: view it as segments of raw values (thread g takes [100 g, 100 g + 100))
        const unsigned m = s_cnt[lane];
        if (s_packed[lane]) {
            const unsigned b = (unsigned)(g * kM2Seg);
            c_own = m <= (unsigned)kMsCap ? (m > b ? min(m - b, (unsigned)kM2Seg) : 0u) : 0u;
            for (unsigned i = 0; i < c_own; ++i) seg[i] = med_unkey(__float_as_uint(seg[i]));
        } else {
            c_own = min(s_seg[g][lane], (unsigned)kM2Seg);
        }
    }
    __syncthreads(); // s_q, s_seg are dead from here on
    for (int i = tid; i < 2 * kMsPix * 8; i += kMsPix * kM2Warps) f_c[i] = 0u;
    if (g == 0) {
        const unsigned m = s_cnt[lane], eq = s_eqt[lane], r = s_r[lane];
        const bool haslo = s_haslo[lane] != 0u;
        const unsigned klo_key = med_key(s_lo[lane]);
        unsigned mode, klo = 0u, khi = 0u, rr = 0u;
        if (!active) {
            mode = 2u;
        } else if (m > (unsigned)kMsCap) {
            mode = 3u;
        } else if (haslo && r < eq) { // the wanted value is lo itself
            klo = klo_key;
            khi = klo_key;
            mode = 2u;
            if (even && !(r + 1 < eq)) { // ... and its successor the smallest candidate, or `above`
                if (m == 0u) khi = s_abv[lane];
                else mode = 1u;
            }
        } else {
            rr = r - (haslo ? eq : 0u);
            mode = 0u;
            if (rr >= m) { mode = 2u; } // only with NaNs in the window (the flag is set): any value will do
        }
        f_mode[lane] = mode;
        if (mode != 3u) { f_klo[lane] = klo; f_khi[lane] = khi; s_r[lane] = rr; f_eq[lane] = 0u; f_m2[lane] = 0u; }
    }
    __syncthreads();
    unsigned seed = (unsigned)(px0 + lane) * 0x9E3779B9u + (unsigned)g * 0x85EBCA6Bu + 0x7F4A7C15u;
    for (int round = 0; round < 64; ++round) { // random pivots shrink a list ~5x per round: 64 is never reached; a guard against a hang, not a limit
        const unsigned mode = f_mode[lane];
        const bool big = mode < 2u && s_cnt[lane] > 32u;
        if (!__syncthreads_or(big ? 1 : 0)) break;
        unsigned *const cnt_now = f_c + (round & 1) * kMsPix * 8 + lane * 8;
        // R1: pivot offers.  Round 0: 4 pivots (the even warps offer), each the sample of the thread's sorted prefix at the
        // wanted rank's quantile -- 4 estimates of the wanted value, close together, so the bin between two of them is small
        // (~70 of ~500 candidates) at half the compares of 8 pivots.  Later rounds: 8 pivots, one random candidate per thread.
        const int np = round == 0 ? 4 : 8;
        seed = seed * 1664525u + 1013904223u;
        {
            float offer = __int_as_float(0x7fc00000);
            if (big && c_own > 0u) {
                unsigned idx = (seed >> 8) % c_own;
                if (round == 0 && ks > 0u && !s_packed[lane]) {
                    const float phi = __fdividef((float)(mode == 1u ? 0u : s_r[lane]) + 0.5f, (float)s_cnt[lane]);
                    idx = min(min(ks, c_own) - 1u, (unsigned)(phi * (float)ks));
                }
                offer = seg[idx];
            }
            f_pv[g * kMsPix + lane] = offer;
        }
        __syncthreads();
        // R2: counts of this segment against the spaxel's sorted pivots (a compare on the ALU pipe + a predicated FADD each)
        float pv[8];
        if (big) {
            float cj[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) { cj[j] = 0.0f; pv[j] = kInf; }
            if (round == 0) {
                float p4[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    p4[j] = f_pv[2 * j * kMsPix + lane];
                    if (p4[j] != p4[j]) p4[j] = kInf; // no offer: sorts last, counts every finite candidate
                }
                m2_sort_regs<4>(p4);
#pragma unroll
                for (int j = 0; j < 4; ++j) pv[j] = p4[j];
#pragma unroll 4
                for (unsigned i = 0; i < c_own; ++i) {
                    const float x = seg[i];
#pragma unroll
                    for (int j = 0; j < 4; ++j)
                        asm("{\n\t.reg .pred p;\n\tsetp.lt.f32 p, %1, %2;\n\t@p add.f32 %0, %0, 0f3F800000;\n\t}" : "+f"(cj[j]) : "f"(x), "f"(pv[j]));
                }
            } else {
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    pv[j] = f_pv[j * kMsPix + lane];
                    if (pv[j] != pv[j]) pv[j] = kInf;
                }
                m2_sort_regs<8>(pv);
#pragma unroll 4
                for (unsigned i = 0; i < c_own; ++i) {
                    const float x = seg[i];
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        asm("{\n\t.reg .pred p;\n\tsetp.lt.f32 p, %1, %2;\n\t@p add.f32 %0, %0, 0f3F800000;\n\t}" : "+f"(cj[j]) : "f"(x), "f"(pv[j]));
                }
            }
#pragma unroll
            for (int j = 0; j < 8; ++j)
                if (cj[j] != 0.0f) atomicAdd(cnt_now + j, (unsigned)cj[j]);
        }
        __syncthreads();
        // R3: the bin that holds the wanted rank; compact this segment to it.  An open side is a NaN bound: `x >= NaN` is
        // false and `x >u NaN` (greater or unordered) is true, so the loop has no flags.
        unsigned below_b = 0u;
        float lo_b = __int_as_float(0x7fc00000), hi_b = __int_as_float(0x7fc00000);
        if (big) {
            const unsigned r = mode == 1u ? 0u : s_r[lane];
            unsigned C[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) C[j] = cnt_now[j];
#pragma unroll
            for (int j = 0; j < 8; ++j)
                if (j < np && C[j] <= r) { below_b = C[j]; lo_b = pv[j]; }
#pragma unroll
            for (int j = 7; j >= 0; --j)
                if (j < np && C[j] > r) hi_b = pv[j];
            float eq_f = 0.0f, ge_f = 0.0f, hmin = kInf;
            const unsigned w0 = (unsigned)__cvta_generic_to_shared(seg);
            unsigned wq_ = w0;
#pragma unroll 4
            for (unsigned i = 0; i < c_own; ++i) {
                const float x = seg[i];
                asm volatile(
                    "{\n\t"
                    ".reg .pred pe, pg, pk;\n\t"
                    "setp.eq.f32 pe, %4, %5;\n\t"
                    "@pe add.f32 %0, %0, 0f3F800000;\n\t"
                    "setp.ge.f32 pg, %4, %6;\n\t"
                    "@pg add.f32 %1, %1, 0f3F800000;\n\t"
                    "@pg min.f32 %2, %2, %4;\n\t"
                    "setp.gtu.and.f32 pk, %4, %5, !pg;\n\t"
                    "@pk st.shared.f32 [%3], %4;\n\t" // at or below slot i: in place
                    "@pk add.u32 %3, %3, 4;\n\t"
                    "}"
                    : "+f"(eq_f), "+f"(ge_f), "+f"(hmin), "+r"(wq_)
                    : "f"(x), "f"(lo_b), "f"(hi_b)
                    : "memory");
            }
            c_own = (wq_ - w0) >> 2;
            if (eq_f != 0.0f) atomicAdd(&f_eq[lane], (unsigned)eq_f);
            if (c_own) atomicAdd(&f_m2[lane], c_own);
            if (ge_f != 0.0f) atomicMin(&s_abv[lane], med_key(hmin)); // candidates >= hi sort below every earlier `above`
        }
        __syncthreads();
        // R4: lane = spaxel: new rank, or the answer when the rank fell on the lower pivot
        if (g == 0 && big) {
            const unsigned eqt = f_eq[lane], m2 = f_m2[lane];
            const unsigned r = mode == 1u ? 0u : s_r[lane];
            const unsigned rr = r - below_b;
            const unsigned pk = med_key(lo_b);
            const bool has_lo_b = lo_b == lo_b;
            if (has_lo_b && rr < eqt) {
                if (mode == 1u) {
                    f_khi[lane] = pk;
                    f_mode[lane] = 2u;
                } else {
                    f_klo[lane] = pk;
                    f_khi[lane] = pk;
                    f_mode[lane] = 2u;
                    if (even && !(rr + 1 < eqt)) {
                        if (m2 == 0u) f_khi[lane] = s_abv[lane];
                        else { f_mode[lane] = 1u; s_r[lane] = 0u; }
                    }
                }
            } else {
                s_r[lane] = rr - (has_lo_b ? eqt : 0u);
            }
            s_cnt[lane] = m2;
            f_eq[lane] = 0u;
            f_m2[lane] = 0u;
#pragma unroll
            for (int j = 0; j < 8; ++j) cnt_now[j] = 0u; // next used two rounds from now
        }
        __syncthreads();
    }
    // at most 32 candidates per live spaxel: gather them at the front of the row and rank them by counting, still in the
    // (warp g, lane = spaxel) layout: thread g ranks the candidates g, g + 8, g + 16, g + 24 of its spaxel against all of them
    f_k[g * kMsPix + lane] = c_own;
    __syncthreads();
    {
        const unsigned mode = f_mode[lane];
        if (mode < 2u && c_own > 0u && g > 0) {
            unsigned off = 0u;
            for (int w = 0; w < g; ++w) off += f_k[w * kMsPix + lane];
            // segment 0 starts at the front and stays; the others land behind its <= 32 candidates
            float *row = reinterpret_cast<float *>(s_list) + lane * kMsRow;
            for (unsigned i = 0; i < c_own; ++i) row[off + i] = seg[i];
        }
        if (g == 0 && mode == 0u && even && s_r[lane] + 1u >= s_cnt[lane]) f_khi[lane] = s_abv[lane]; // the successor is beyond S
    }
    __syncthreads();
    {
        const unsigned mode = f_mode[lane];
        const unsigned m = s_cnt[lane];
        if (mode < 2u && m > 0u) {
            const float *row = reinterpret_cast<const float *>(s_list) + lane * kMsRow;
            const unsigned r = mode == 1u ? 0u : s_r[lane];
            float xc[4], lt[4], le[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const unsigned i = (unsigned)g + 8u * k;
                xc[k] = row[i < m ? i : 0u];
                lt[k] = 0.0f;
                le[k] = 0.0f;
            }
            for (unsigned j = 0; j < m; ++j) {
                const float y = row[j];
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    asm("{\n\t.reg .pred p;\n\tsetp.lt.f32 p, %1, %2;\n\t@p add.f32 %0, %0, 0f3F800000;\n\t}" : "+f"(lt[k]) : "f"(y), "f"(xc[k]));
                    asm("{\n\t.reg .pred p;\n\tsetp.le.f32 p, %1, %2;\n\t@p add.f32 %0, %0, 0f3F800000;\n\t}" : "+f"(le[k]) : "f"(y), "f"(xc[k]));
                }
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                if ((unsigned)g + 8u * k >= m) continue;
                const unsigned l = (unsigned)lt[k], h = (unsigned)le[k];
                const unsigned key = med_key(xc[k]);
                // equal candidates write the same value (a zero of either sign: median_zero_fix_kernel decides)
                if (mode == 1u) {
                    if (l == 0u) f_khi[lane] = key;
                } else {
                    if (l <= r && r < h) f_klo[lane] = key;
                    if (l <= r + 1u && r + 1u < h) f_khi[lane] = key;
                }
            }
        }
    }
    __syncthreads();
    if (g == 0 && active && f_mode[lane] != 3u) a.out[mypx] = med_finish(f_klo[lane], f_khi[lane], even);
    for (int p = g; p < kMsPix; p += kM2Warps) {
        const int64_t px = px0 + p;
        if (px >= a.n2) continue; // warp-uniform
        const unsigned mode = f_mode[p];
        if (mode != 3u) continue;
        unsigned *list = s_list + p * kMsRow;
        unsigned klo = 0u, khi = 0u;
        {
            // the re-collected bin did not fit the list (heavy ties next to the median, or a tied bracket): refine it against
            // the global column until it does, then the warp-wide rounds
            unsigned m = s_cnt[p];
            const float flo = s_lo[p], fhi = s_hi[p];
            bool bhas_lo = s_haslo[p] != 0u;
            // bounds as keys, consistent with the float compares: values > (+-)0 are the keys above +0.0, values < (+-)0 the keys below -0.0
            unsigned blo = bhas_lo ? med_key(flo == 0.0f ? 0.0f : flo) : 0u;
            unsigned bhi = s_hashi[p] != 0u ? med_key(fhi == 0.0f ? -0.0f : fhi) : kFull;
            const unsigned eq = s_eqt[p];
            unsigned above = s_abv[p];
            unsigned r = s_r[p];
            const float *pcol = a.cube + a.start * a.n2 + px;
            bool done = false;
            unsigned sd = (unsigned)px * 0x9E3779B9u + 0x7F4A7C15u;
            if (bhas_lo && r < eq) { // the wanted value is lo itself
                klo = blo;
                khi = blo;
                if (even && !(r + 1 < eq)) {
                    unsigned mn = kFull;
                    for (int k = lane; k < n; k += 32) {
                        const unsigned key = ms_key(ms_ld(pcol + (size_t)k * a.n2));
                        if (key > blo && key < bhi) mn = min(mn, key);
                    }
                    mn = __reduce_min_sync(kFull, mn);
                    khi = mn != kFull ? mn : above;
                }
                done = true;
            } else {
                if (bhas_lo) r -= eq;
                while (!done && m > (unsigned)kMsCap) {
                    done = ms_refine(pcol, a.n2, n, list, blo, bhi, bhas_lo, r, m, above, even, lane, sd, klo, khi);
                    sd = sd * 1664525u + 1013904223u;
                }
            }
            if (!done) {
                MedState st{(int)m, r, above, above != kFull};
                while (!done && st.m > 32) {
                    done = med_round(list, st, even, lane, sd, klo, khi);
                    sd = sd * 1664525u + 1013904223u;
                }
                if (!done) med_rank32(list, st, even, lane, klo, khi);
            }
        }
        if (lane == 0) a.out[px] = med_finish(klo, khi, even);
    }
    if (a.trace) { __syncthreads(); ms_trace(a, 4, t_tr); }
}

// ---------------------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------------------
static size_t a0_state_bytes(int op, int mode) {
    switch (op) {
    case A0_VARIANCE: case A0_STDDEV: return mode == SUM_SEQUENTIAL ? sizeof(VarSeq::State) : sizeof(VarComp::State);
    case A0_ARGFOLD: return sizeof(ArgFold::State);
    case A0_CENTROID: return mode == SUM_SEQUENTIAL ? sizeof(CentroidSeq::State) : sizeof(CentroidComp::State);
    }
    return 0;
}

Axis0Plan axis0_plan(const Axis0Args &a, int sm_count) {
    Axis0Plan pl{};
    const bool aligned = (a.n2 % 4 == 0) && ((reinterpret_cast<uintptr_t>(a.cube) & 15u) == 0);
    pl.vec = aligned ? 4 : 1;
    pl.block = 256;
    const int64_t n = a.end - a.start;
    const bool mergeable = a.op == A0_ARGFOLD || a.mode != SUM_SEQUENTIAL;
    // An order-dependent reduction cannot split the spectral axis: on a small image the only parallelism is the pixels,
    // so give every pixel its own thread and spread the warps over the SMs (latency-bound chain per pixel either way)
    if (!mergeable && (a.n2 + 256 * pl.vec - 1) / (256 * pl.vec) < sm_count) {
        pl.vec = 1;
        pl.block = a.n2 >= (int64_t)64 * sm_count ? 64 : 32;
    }
    pl.ntiles = (a.n2 + (int64_t)pl.block * pl.vec - 1) / ((int64_t)pl.block * pl.vec);
    pl.ksplit = 1;
    const int64_t want = 2 * (int64_t)sm_count;
    if (mergeable && pl.ntiles < want && n >= 32) {
        int64_t ks = (want + pl.ntiles - 1) / pl.ntiles;
        const int64_t max_ks = n / 16; // at least 16 frames per chunk
        if (ks > max_ks) ks = max_ks;
        if (ks > 1024) ks = 1024;
        if (ks > 1) pl.ksplit = (int)ks;
    }
    pl.chunk = (n + pl.ksplit - 1) / pl.ksplit;
    pl.ksplit = (int)((n + pl.chunk - 1) / pl.chunk);
    pl.partial_bytes = pl.ksplit > 1 ? (size_t)pl.ksplit * (size_t)a.n2 * a0_state_bytes(a.op, a.mode) : 0;
    return pl;
}

template <class F, bool WAVE> static void a0_go(const Axis0K &k, const Axis0Plan &pl, cudaStream_t s) {
    const dim3 grid((unsigned)pl.ntiles, (unsigned)pl.ksplit);
    if (pl.vec == 4) axis0_kernel<F, 4, 8, WAVE><<<grid, pl.block, 0, s>>>(k);
    else if (!F::kMergeable && pl.block <= 64) { // small image, order-dependent
        Axis0K kk = k;
        const int64_t n = k.end - k.start;
        kk.wk_table = WAVE && n * (int64_t)sizeof(float) <= 40 * 1024;
        axis0_kernel<F, 1, 16, WAVE, true><<<grid, pl.block, kk.wk_table ? (size_t)n * sizeof(float) : 0, s>>>(kk);
    }
    else axis0_kernel<F, 1, 16, WAVE><<<grid, pl.block, 0, s>>>(k);
    count_launch();
    if (pl.ksplit > 1) {
        axis0_merge_kernel<F><<<(unsigned)((k.n2 + 255) / 256), 256, 0, s>>>(k);
        count_launch();
    }
}

cudaError_t axis0_launch(const Axis0Args &a, const Axis0Plan &pl, void *partials, unsigned *flag, cudaStream_t s) {
    if (a.n2 <= 0) return cudaSuccess;
    if (a.end <= a.start) return cudaErrorInvalidValue;
    if (pl.ksplit > 1 && partials == nullptr) return cudaErrorInvalidValue;
    Axis0K k{};
    k.cube = a.cube; k.n2 = a.n2; k.nz = a.nz; k.start = a.start; k.end = a.end;
    k.chunk = pl.chunk; k.ksplit = pl.ksplit;
    k.is_min = a.is_min; k.r1 = a.r1; k.r2 = a.r2; k.out2_kind = a.out2_kind;
    k.sqrt_out = a.op == A0_STDDEV;
    k.use_wcs = a.use_wcs; k.crval = a.crval; k.crpix = a.crpix; k.cdelt = a.cdelt;
    k.mask_lo = a.mask_lo; k.mask_hi = a.mask_hi;
    k.out = a.out; k.out2 = a.out2; k.partials = partials; k.flag = flag;
    const bool seq = a.mode == SUM_SEQUENTIAL;
    switch (a.op) {
    case A0_VARIANCE: case A0_STDDEV:
        if (seq) a0_go<VarSeq, false>(k, pl, s); else a0_go<VarComp, false>(k, pl, s);
        break;
    case A0_ARGFOLD: a0_go<ArgFold, false>(k, pl, s); break;
    case A0_CENTROID:
        if (seq) a0_go<CentroidSeq, true>(k, pl, s); else a0_go<CentroidComp, true>(k, pl, s);
        break;
    case A0_CENTROID_MASK:
        if (flag == nullptr || a.mask_lo == nullptr || a.mask_hi == nullptr) return cudaErrorInvalidValue;
        if (seq) centroid_mask_kernel<false><<<(unsigned)((a.n2 + 255) / 256), 256, 0, s>>>(k);
        else centroid_mask_kernel<true><<<(unsigned)((a.n2 + 255) / 256), 256, 0, s>>>(k);
        count_launch();
        break;
    default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

// shared memory for staged keys: 227 KB opt-in per CTA; prefer a tile that lets two CTAs share an SM, so one CTA's
// staging overlaps the other's select
static constexpr size_t kMedianSmemMax = 224 * 1024;

static int median_npad(int64_t n, int P) {
    const int want = P >= 32 ? 1 : 32 / P; // npad % 32 == 32/P: a warp's staging stores (P spaxels x 32/P frames) hit 32 banks
    int64_t np = n;
    while ((np % 32) != (want % 32)) ++np;
    return (int)np;
}
static size_t median_select_smem(int64_t n, int P) {
    return (size_t)P * median_npad(n, P) * sizeof(unsigned) + (size_t)(8 * 256 + 8 * 16) * sizeof(unsigned);
}

int median_pixels_per_cta(int64_t n2, int64_t n) {
    int fits_once = 0;
    for (int P = 32; P >= 1; P >>= 1) {
        const size_t bytes = median_select_smem(n, P);
        if (bytes <= kMedianSmemMax && fits_once == 0) fits_once = P;
        if (bytes * 2 <= kMedianSmemMax + 3 * 1024) { // two resident CTAs
            int best = P;
            while (best > 4 && (n2 + best - 1) / best < 2 * 148) best >>= 1; // small images: more, narrower CTAs
            return best;
        }
    }
    return fits_once; // 0: not even one spaxel fits -> streaming kernel
}

// The reference sorts with `sort_by(partial_cmp)`: a STABLE sort under which -0.0 == +0.0, so zeros keep their frame order,
// whereas the select kernels order keys as integers (-0.0 before +0.0).  Only the sign of a zero result can differ, and only
// when the window holds zeros of both signs.  This pass visits the pixels whose result is zero: with L = #(v < 0) the zeros
// occupy the sorted ranks L, L+1, ... in frame order, so the middle element(s) are the (mid - L)-th zero(s) of the window.
__global__ void __launch_bounds__(256) median_zero_fix_kernel(const float *__restrict__ cube, int64_t n2, int64_t start, int64_t n,
                                                               float *__restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n2) return;
    if (out[i] != 0.0f) return; // NaN and non-zero results are already final
    const float *col = cube + start * n2 + i;
    int64_t less = 0, zeros = 0;
    for (int64_t k = 0; k < n; ++k) {
        const float v = col[k * n2];
        less += v < 0.0f;
        zeros += v == 0.0f;
    }
    const int64_t hi = n / 2 - less, lo = (n % 2 == 1) ? hi : hi - 1; // ranks of the middle element(s) among the zeros
    if (lo < 0 || hi >= zeros) return; // one of the middles is not a zero: x + (+-0) = x, the result is already exact
    float zlo = 0.0f, zhi = 0.0f;
    int64_t seen = 0;
    for (int64_t k = 0; k < n && seen <= hi; ++k) {
        const float v = col[k * n2];
        if (v == 0.0f) {
            if (seen == lo) zlo = v;
            if (seen == hi) zhi = v;
            ++seen;
        }
    }
    out[i] = (n % 2 == 1) ? zhi : __fdiv_rn(__fadd_rn(zhi, zlo), 2.0f);
}

static cudaError_t median_dispatch(const float *cube, int64_t n2, int64_t start, int64_t end, float *out, unsigned *flag, cudaStream_t s);

cudaError_t median_launch(const float *cube, int64_t n2, int64_t start, int64_t end, float *out, unsigned *flag, cudaStream_t s) {
    if (n2 <= 0) return cudaSuccess;
    cudaError_t e = median_dispatch(cube, n2, start, end, out, flag, s);
    if (e != cudaSuccess) return e;
    median_zero_fix_kernel<<<(unsigned)((n2 + 255) / 256), 256, 0, s>>>(cube, n2, start, end - start, out);
    count_launch();
    return cudaGetLastError();
}

static cudaError_t median_dispatch(const float *cube, int64_t n2, int64_t start, int64_t end, float *out, unsigned *flag, cudaStream_t s) {
    if (n2 <= 0) return cudaSuccess;
    const int64_t n = end - start;
    if (n < 1 || n > 0x7FFFFFFF || flag == nullptr) return cudaErrorInvalidValue;
    MedianK k{cube, n2, start, n, 0, 0, out, flag, nullptr};
    if (n <= kMedRankMaxN) {
        const unsigned grid = (unsigned)((n2 + kMedSortPixels - 1) / kMedSortPixels);
        if (n <= 8) median_sort_kernel<8><<<grid, kMedSortPixels, 0, s>>>(k);
        else if (n <= 16) median_sort_kernel<16><<<grid, kMedSortPixels, 0, s>>>(k);
        else if (n <= 32) median_sort_kernel<32><<<grid, kMedSortPixels, 0, s>>>(k);
        else if (n <= 64) median_sort_kernel<64><<<grid, kMedSortPixels, 0, s>>>(k);
        else median_sort_kernel<128><<<grid, kMedSortPixels, 0, s>>>(k);
        count_launch();
        return cudaGetLastError();
    }
    // Long windows on images with enough 32-spaxel tiles to fill the GPU are streamed (two coalesced passes, pivots
    // prune 15/16 of the keys); so is any window too long for shared memory.  Short windows and small images keep the
    // shared-memory select (profiles/r1_median_phases.txt).
    const int P = n > (int64_t)(kMedianSmemMax / sizeof(unsigned)) ? 0 : median_pixels_per_cta(n2, n);
    static const int64_t stream_min = getenv("AFK_MEDIAN_STREAM_MIN") ? atoll(getenv("AFK_MEDIAN_STREAM_MIN")) : kMedStreamMinFrames; // tuning knob
    if (P == 0 || (n >= stream_min && n2 >= kMedStreamMinPixels)) {
        constexpr size_t smem_ms = (size_t)kMsPix * kMsRow * sizeof(unsigned);
        static bool attr_ms[16] = {};
        int dev_ms = 0;
        cudaGetDevice(&dev_ms);
        if (dev_ms >= 0 && dev_ms < 16 && !attr_ms[dev_ms]) {
            cudaError_t e = cudaFuncSetAttribute(median_stream_kernel<7, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_ms);
            if (e == cudaSuccess) e = cudaFuncSetAttribute(median_stream_kernel<15, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_ms);
            if (e == cudaSuccess) e = cudaFuncSetAttribute(median_stream1_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_ms);
#define AFK_M2_ATTR(SS_) if (e == cudaSuccess) e = cudaFuncSetAttribute(median_stream2_kernel<SS_>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_ms)
            AFK_M2_ATTR(2); AFK_M2_ATTR(3); AFK_M2_ATTR(4); AFK_M2_ATTR(5); AFK_M2_ATTR(6); AFK_M2_ATTR(7); AFK_M2_ATTR(8); AFK_M2_ATTR(9);
#undef AFK_M2_ATTR
            if (e != cudaSuccess) return e;
            attr_ms[dev_ms] = true;
        }
        const unsigned grid_ms = (unsigned)((n2 + kMsPix - 1) / kMsPix);
        static const int single_pass = getenv("AFK_MEDIAN_SINGLE_PASS") ? atoi(getenv("AFK_MEDIAN_SINGLE_PASS")) : 1; // 0: the two-pass kernel (A/B); 2: from 801 frames
        static const int64_t one_pass_min = getenv("AFK_MEDIAN_ONEPASS_MIN") ? atoll(getenv("AFK_MEDIAN_ONEPASS_MIN")) : kMsCap + 1;
        // Measured on the ALMA cube (profiles/r2_median_one_pass_v2.json): the rebuilt one-pass kernel beats the two-pass kernel on
        // every window its 800-slot rows allow (1025 frames: 3.05 vs 3.48 ms, 2000: 3.81 vs 5.39, 3840: 6.0 vs 9.1) -> used from
        // 801 frames up (AFK_MEDIAN_ONEPASS_MIN moves the switch, AFK_MEDIAN_SINGLE_PASS=0 selects the two-pass kernel).
        if (single_pass && (n >= one_pass_min || single_pass == 2) && n > kMsCap && n <= kMs1MaxFrames) {
            static const int trace = getenv("AFK_MEDIAN_TRACE") ? atoi(getenv("AFK_MEDIAN_TRACE")) : 0;
            // AFK_MEDIAN_ONEPASS_V=1: the first one-pass kernel (shared atomics, 16 warps) for A/B
            static const int version = getenv("AFK_MEDIAN_ONEPASS_V") ? atoi(getenv("AFK_MEDIAN_ONEPASS_V")) : 2;
            // half-widths of the bracket and of the wing in quantile units: z sigma of the sample quantile near the median,
            // sigma = 1/2 sqrt((1 - M/n) / M) with M ~ n / ceil(n / 512) samples per spaxel (AFK_MEDIAN_Z, default 3; wing 2 z + 1)
            static const double z = getenv("AFK_MEDIAN_Z") ? atof(getenv("AFK_MEDIAN_Z")) : 3.0;
            const int64_t mine = (n + kM2Warps - 1) / kM2Warps, ssamp = (mine + kM2Smp - 1) / kM2Smp;
            const double msamp = (double)kM2Warps * (double)((mine + ssamp - 1) / ssamp);
            const double sigma_q = 0.5 * std::sqrt(std::max(0.0, 1.0 - msamp / (double)n) / msamp) + 0.25 / msamp;
            const float wq = (float)std::min(0.2, z * sigma_q), wq2 = (float)std::min(0.45, (2.0 * z + 1.0) * sigma_q);
            const unsigned block = version == 1 ? kMsPix * kMsWarps : kMsPix * kM2Warps;
            MedianK kt = k;
            unsigned long long *d = nullptr, h[8] = {};
            if (trace) { // diagnostics: synchronises
                cudaMalloc(&d, sizeof h);
                cudaMemset(d, 0, sizeof h);
                kt.trace = d;
            }
            if (version == 1) median_stream1_kernel<<<grid_ms, block, smem_ms, s>>>(kt);
            else {
                // Thread-block clusters for DRAM locality, not for shared memory: a CTA reads 128 contiguous bytes of a frame at a
                // time; the CTAs of a cluster start together on neighbouring SMs and walk adjacent tiles in step, so the memory
                // system sees 128 B x cluster-size contiguous requests close in time (AFK_MEDIAN_CLUSTER = 1|2|4|8)
                static const int cl = getenv("AFK_MEDIAN_CLUSTER") ? atoi(getenv("AFK_MEDIAN_CLUSTER")) : 2;
                const unsigned csize = (cl == 2 || cl == 4 || cl == 8) ? (unsigned)cl : 1u;
                cudaLaunchConfig_t cfg = {};
                cfg.gridDim = dim3((grid_ms + csize - 1) / csize * csize);
                cfg.blockDim = dim3(block);
                cfg.dynamicSmemBytes = smem_ms;
                cfg.stream = s;
                cudaLaunchAttribute at[1];
                at[0].id = cudaLaunchAttributeClusterDimension;
                at[0].val.clusterDim.x = csize;
                at[0].val.clusterDim.y = 1;
                at[0].val.clusterDim.z = 1;
                cfg.attrs = at;
                cfg.numAttrs = 1;
                cudaError_t le = cudaErrorInvalidValue;
                switch ((int)ssamp) { // the sample stride is a template parameter: whole blocks of SS - 1 streamed frames per batch
#define AFK_M2_GO(SS_) case SS_: le = cudaLaunchKernelEx(&cfg, median_stream2_kernel<SS_>, kt, wq, wq2); break
                    AFK_M2_GO(2); AFK_M2_GO(3); AFK_M2_GO(4); AFK_M2_GO(5); AFK_M2_GO(6); AFK_M2_GO(7); AFK_M2_GO(8); AFK_M2_GO(9);
#undef AFK_M2_GO
                }
                if (le != cudaSuccess) return le;
            }
            if (trace) {
                cudaStreamSynchronize(s);
                cudaMemcpy(h, d, sizeof h, cudaMemcpyDeviceToHost);
                cudaFree(d);
                const double tot = (double)(h[0] + h[1] + h[2] + h[3] + h[4]);
                fprintf(stderr, "[afk median trace] one-pass v%d n=%lld CTAs=%u cycles/CTA: sample loads %.0f, sort %.0f, stream %.0f, miss re-collect %.0f, finish %.0f (%.1f%% %.1f%% %.1f%% %.1f%% %.1f%%)\n",
                        version, (long long)n, grid_ms, h[0] / (double)grid_ms, h[1] / (double)grid_ms, h[2] / (double)grid_ms, h[3] / (double)grid_ms, h[4] / (double)grid_ms,
                        100 * h[0] / tot, 100 * h[1] / tot, 100 * h[2] / tot, 100 * h[3] / tot, 100 * h[4] / tot);
                fprintf(stderr, "[afk median trace] spaxels %lld: bracket missed %llu (%.2f%%), list overflowed %llu, mean keys in the bracket %.1f\n", (long long)n2,
                        h[5], 100.0 * h[5] / (double)n2, h[7], h[6] / (double)n2);
            }
        }
        else if (n <= kMedStreamWideBins) median_stream_kernel<7, 4><<<grid_ms, kMsPix * kMsWarps, smem_ms, s>>>(k);
        else median_stream_kernel<15, 2><<<grid_ms, kMsPix * kMsWarps, smem_ms, s>>>(k);
        count_launch();
        return cudaGetLastError();
    }
    k.P = P;
    k.npad = median_npad(n, P);
    const size_t smem = median_select_smem(n, P);
    static bool attr_set[16] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 16 && !attr_set[dev]) {
        cudaError_t e = cudaFuncSetAttribute(median_select_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMedianSmemMax);
        if (e != cudaSuccess) return e;
        attr_set[dev] = true;
    }
    median_select_kernel<<<(unsigned)((n2 + P - 1) / P), 256, smem, s>>>(k);
    count_launch();
    return cudaGetLastError();
}

} // namespace afk
