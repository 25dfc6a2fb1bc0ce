// stats.cu -- display statistics aflak_plot computes for every new image beside the colour mapping:
// the 100-bin (256-bin per channel for colour images) histogram, the normalised median and median absolute deviation,
// and the strided sample zscale works on.
//
// Replaces aflak_plot/src/imshow/hist.rs:10-66, lims.rs:37-106 (two full sorts of the image on the UI thread) and the
// sampling loop of zscale.rs:101-126.  All results are exact: bin indices use the reference's f32 step sequence, the
// medians are order statistics found by a 4-pass radix select over order-preserving keys of the values the reference
// would have sorted (computed on the fly, never materialised).  A 4096x4096 image (64 MiB) stays L2-resident over the
// passes; every pass is a streaming read with shared-memory privatised histograms.
#include "kernels.h"

#include <cstdio>
#include <cstdlib>

#include <cstdint>

namespace afk {

__device__ __forceinline__ float4 st_ld(const float4 *p) {
    float4 r;
    asm("ld.global.nc.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}

// ---- hist::histogram / histogram_color ---------------------------------------------------------------------------------
// i = ((val - min) / (max - min) * NB) as usize, clamped to NB-1 (hist.rs:26-31): NaN and negatives land in bin 0
__device__ __forceinline__ unsigned hist_bin(float v, float mn, float range, float nb_f, unsigned nb) {
    const float t = __fmul_rn(__fdiv_rn(__fsub_rn(v, mn), range), nb_f);
    unsigned i;
    if (!(t > 0.0f)) i = 0u;
    else if (t >= 4294967296.0f) i = 0xFFFFFFFFu;
    else i = (unsigned)t;
    return i >= nb ? nb - 1u : i;
}

// A thread meets long runs of the same bin (smooth images, concentrated distributions): count the run in a register
// and touch shared memory only when the bin changes.
struct RunCounter {
    unsigned bin = 0xFFFFFFFFu, run = 0u;
    __device__ __forceinline__ void add(unsigned b, unsigned *hist) {
        if (b == bin) { ++run; return; }
        if (run) atomicAdd(&hist[bin], run);
        bin = b;
        run = 1u;
    }
    __device__ __forceinline__ void flush(unsigned *hist) {
        if (run) atomicAdd(&hist[bin], run);
        run = 0u;
    }
};

// The same counter without branches, for the select passes (issue-bound: the branchy form costs ~25 instructions and four
// branches per key, profiles/r1_stats_ncu_full.json).  `match` = the key takes part; a run that ends is added to its
// bin by a predicated reduction.  Starts at (bin 0, run 0): the first break adds 0 to bin 0.
struct RunCounterPred {
    unsigned bin = 0u, run = 0u;
    __device__ __forceinline__ void add_if(bool match, unsigned b, unsigned hist_saddr) {
        const bool brk = match && b != bin;
        asm volatile(
            "{\n\t"
            ".reg .pred p;\n\t"
            "setp.ne.u32 p, %0, 0;\n\t"
            "@p red.shared.add.u32 [%1], %2;\n\t"
            "}"
            :
            : "r"((unsigned)brk), "r"(hist_saddr + bin * 4u), "r"(run)
            : "memory");
        run = brk ? 1u : (match ? run + 1u : run);
        bin = brk ? b : bin;
    }
    __device__ __forceinline__ void flush(unsigned *hist) {
        if (run) atomicAdd(&hist[bin], run);
        run = 0u;
    }
};

constexpr int kHistWarps = 8;

// counts: [nbins][channels] (bin-major like Vec<[Bin; 3]>), 64-bit
// host_out (may be null, then ticket is unused): mapped pinned host memory for the [nb][channels] counts.  The last CTA of the
// grid (ticket) copies the finished counts there with plain stores and returns counts[] and the ticket to zero, so that the
// same device block serves the next launch on the stream without a memset and the host needs no copy, only a synchronisation.
__global__ void __launch_bounds__(256) histogram_kernel(const float *__restrict__ img, int64_t n_per_channel, int channels, float mn,
                                                         float mx, unsigned nb, unsigned long long *counts, unsigned *ticket,
                                                         unsigned long long *host_out) {
    extern __shared__ unsigned s_hist[]; // [kHistWarps][nb]
    for (unsigned i = threadIdx.x; i < kHistWarps * nb; i += blockDim.x) s_hist[i] = 0u;
    __syncthreads();
    unsigned *my = s_hist + (threadIdx.x >> 5) * nb;
    const float range = __fsub_rn(mx, mn), nb_f = (float)nb;
    const int c = blockIdx.y;
    const float *src = img + (size_t)c * n_per_channel;
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nthreads = (int64_t)gridDim.x * blockDim.x;
    RunCounterPred rc;
    const unsigned my_saddr = (unsigned)__cvta_generic_to_shared(my);
    int64_t done = 0;
    if ((reinterpret_cast<uintptr_t>(src) & 15u) == 0) {
        const int64_t n4 = n_per_channel >> 2;
        for (int64_t i = tid; i < n4; i += nthreads) {
            const float4 v = st_ld(reinterpret_cast<const float4 *>(src) + i);
            rc.add_if(true, hist_bin(v.x, mn, range, nb_f, nb), my_saddr);
            rc.add_if(true, hist_bin(v.y, mn, range, nb_f, nb), my_saddr);
            rc.add_if(true, hist_bin(v.z, mn, range, nb_f, nb), my_saddr);
            rc.add_if(true, hist_bin(v.w, mn, range, nb_f, nb), my_saddr);
        }
        done = n4 << 2;
    }
    for (int64_t i = done + tid; i < n_per_channel; i += nthreads) rc.add_if(true, hist_bin(src[i], mn, range, nb_f, nb), my_saddr);
    rc.flush(my);
    __syncthreads();
    for (unsigned b = threadIdx.x; b < nb; b += blockDim.x) {
        unsigned t = 0u;
#pragma unroll
        for (int w = 0; w < kHistWarps; ++w) t += s_hist[w * nb + b];
        if (t) atomicAdd(&counts[(size_t)b * channels + c], (unsigned long long)t);
    }
    if (host_out) {
        __shared__ unsigned s_last;
        __threadfence();
        __syncthreads();
        if (threadIdx.x == 0) s_last = atomicAdd(ticket, 1u) == gridDim.x * gridDim.y - 1u ? 1u : 0u;
        __syncthreads();
        if (s_last) {
            __threadfence();
            const unsigned total = nb * (unsigned)channels;
            for (unsigned i = threadIdx.x; i < total; i += blockDim.x) {
                *reinterpret_cast<volatile unsigned long long *>(host_out + i) = __ldcg(&counts[i]);
                counts[i] = 0ull;
            }
            if (threadIdx.x == 0) *ticket = 0u;
        }
    }
}

cudaError_t histogram_launch(const float *img, int64_t n_per_channel, int channels, float mn, float mx, int nbins,
                             unsigned long long *counts_dev, int sm_count, cudaStream_t s, unsigned *ticket_dev, unsigned long long *host_out) {
    if ((ticket_dev == nullptr) != (host_out == nullptr)) return cudaErrorInvalidValue;
    if (n_per_channel <= 0) return cudaSuccess;
    if (nbins < 1 || nbins > 1024 || channels < 1 || channels > 3) return cudaErrorInvalidValue;
    int64_t need = ((n_per_channel + 3) / 4 + 255) / 256;
    const int64_t cap = (int64_t)sm_count * 4;
    const dim3 grid((unsigned)(need < 1 ? 1 : (need < cap ? need : cap)), (unsigned)channels);
    histogram_kernel<<<grid, 256, (size_t)kHistWarps * nbins * sizeof(unsigned), s>>>(img, n_per_channel, channels, mn, mx, (unsigned)nbins, counts_dev, ticket_dev, host_out);
    count_launch();
    return cudaGetLastError();
}

// ---- lims::get_vmed_normalized / get_vmad_normalized: exact order statistics by radix select ---------------------------
__device__ __forceinline__ unsigned st_key(float f) {
    const unsigned b = __float_as_uint(f);
    return b ^ ((unsigned)((int)b >> 31) | 0x80000000u); // negative: ~b, else b | sign bit -- one shift and one LOP3
}
__device__ __forceinline__ float st_unkey(unsigned k) {
    const unsigned b = (k & 0x80000000u) ? (k & 0x7FFFFFFFu) : ~k;
    return __uint_as_float(b);
}

// the value the reference pushes into `data` before sorting: d = (v - vmin) / (vmax - vmin) (lims.rs:48), or
// |d - vmed| (lims.rs:87); kept only when neither NaN nor infinite
// p.raw: d is a monotone function of v (two IEEE operations with a positive divisor), so sorting d is sorting v and
// data[i] = d(v_sorted[i]); when vmin, vmax and vmax - vmin are finite and vmax > vmin (checked on the host) every
// non-NaN pixel gives a finite d in [0, 1] and is kept.  The selection then runs on the keys of v -- no division per pixel
// and pass -- and stat_finish_kernel normalises the one or two pixels it ends on.
__device__ __forceinline__ float stat_normalise(float v, const StatParams &p) {
    return __fdiv_rn(__fsub_rn(v, p.vmin), __fsub_rn(p.vmax, p.vmin));
}
__device__ __forceinline__ bool stat_value(float v, const StatParams &p, float vmed, unsigned *key) {
    if (p.raw) {
        if (v != v) return false;
        *key = st_key(v);
        return true;
    }
    float d = stat_normalise(v, p);
    if (p.use_vmed) d = fabsf(__fsub_rn(d, vmed));
    if (!(fabsf(d) < __int_as_float(0x7f800000))) return false; // NaN or +-Inf
    *key = st_key(d);
    return true;
}

// Radix select, 8 bits per pass, most significant digit first.  Three passes read the image (digits 31..24, 23..16,
// 15..8); the third also copies the keys that share the 16-bit prefix found so far -- about 1 % of a typical image -- to a
// candidate buffer and tracks the smallest key above that bucket, so the last digit, the even-count successor and the final
// value come from one small single-CTA kernel over the candidates instead of two more passes over the image.  The bin
// pick that follows each histogram is done by the last CTA to finish (ticket), not by a separate launch.
// 4 launches and 3 image passes per statistic (were 10 and 5).

// one warp: find the bin holding rank kth among the candidates, narrow the prefix, clear the histogram for the next pass
template <bool GLOBAL> __device__ __forceinline__ void stat_pick(StatState *state, unsigned *hist, int shift, int lane) {
    if (state->error) return;
    unsigned long long c[8], tot = 0ull;
#pragma unroll
    for (int j = 0; j < 8; ++j) { c[j] = GLOBAL ? __ldcg(&hist[lane * 8 + j]) : hist[lane * 8 + j]; tot += c[j]; }
    unsigned long long incl = tot;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned long long t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    const unsigned long long total = __shfl_sync(0xffffffffu, incl, 31);
    unsigned long long kth = state->kth;
    if (shift == 24) { // first pass: the candidates are all the kept values
        if (total == 0ull) {
            __syncwarp();
            if (lane == 0) state->error = 1; // "Empty data!" (lims.rs:53, 92)
            return;
        }
        const int even = (total % 2ull) == 0ull;
        kth = even ? total / 2ull - 1ull : total / 2ull; // data[mid - 1] (then data[mid] via the successor) or data[mid]
        if (lane == 0) { state->n = total; state->even = even; }
    }
    unsigned long long excl = incl - tot;
    int digit = -1;
    unsigned long long below = 0ull, cnt = 0ull;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        if (digit < 0 && kth < excl + c[j]) { digit = lane * 8 + j; below = excl; cnt = c[j]; }
        excl += c[j];
    }
    const unsigned ballot = __ballot_sync(0xffffffffu, digit >= 0);
    const int src = __ffs(ballot) - 1;
    digit = __shfl_sync(0xffffffffu, digit, src);
    below = __shfl_sync(0xffffffffu, below, src);
    cnt = __shfl_sync(0xffffffffu, cnt, src);
#pragma unroll
    for (int j = 0; j < 8; ++j) hist[lane * 8 + j] = 0u;
    __syncwarp(); // every lane has read the old state
    if (lane == 0) {
        state->kth = kth - below;
        state->eq_count = cnt;
        state->prefix |= (unsigned)digit << shift;
        state->mask |= 255u << shift;
    }
}

constexpr int kStatCandSmem = 4096; // candidates a CTA stages before one reservation in the global buffer

// SHIFT = 24, 16: histogram of the next digit among the keys that match the prefix.  SHIFT = 8: also the candidate copy.
template <int SHIFT, bool RAW>
__global__ void __launch_bounds__(256) stat_pass_kernel(const float *__restrict__ img, int64_t n, StatParams p, const StatState *prev,
                                                         StatState *state, unsigned *hist, unsigned *cand, unsigned cand_cap) {
    __shared__ unsigned s_hist[kHistWarps][256];
    __shared__ unsigned s_cand[SHIFT == 8 ? kStatCandSmem : 1];
    __shared__ unsigned s_ccount, s_base, s_last;
    for (int i = threadIdx.x; i < kHistWarps * 256; i += blockDim.x) (&s_hist[0][0])[i] = 0u;
    if (threadIdx.x == 0) s_ccount = 0u;
    __syncthreads();
    // state->error only changes in the pick of the first pass, which every CTA of the later passes sees complete
    if (SHIFT != 24 && state->error) return;
    unsigned *my = s_hist[threadIdx.x >> 5];
    const unsigned prefix = state->prefix, mask = state->mask;
    const float vmed = p.use_vmed ? prev->result : 0.0f;
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nthreads = (int64_t)gridDim.x * blockDim.x;
    RunCounterPred rc;
    const unsigned my_saddr = (unsigned)__cvta_generic_to_shared(my);
    unsigned above = 0xFFFFFFFFu;
    auto one = [&](float v) {
        unsigned key = 0u;
        bool kept;
        if (RAW) { // compile-time copy of stat_value's raw branch
            kept = !(v != v);
            key = st_key(v);
        } else {
            kept = stat_value(v, p, vmed, &key);
        }
        const bool match = kept && (key & mask) == prefix;
        rc.add_if(match, (key >> SHIFT) & 255u, my_saddr);
        if (SHIFT == 8) {
            if (match) { // ~1 % of the keys
                const unsigned slot = atomicAdd(&s_ccount, 1u);
                if (slot < (unsigned)kStatCandSmem) {
                    s_cand[slot] = key;
                } else { // a CTA with more candidates than it can stage (long runs of equal values): straight to the buffer
                    const unsigned at = atomicAdd(&state->cand_count, 1u);
                    if (at < cand_cap) cand[at] = key;
                }
            } else if (kept && key > (prefix | 0x0000FFFFu)) {
                above = min(above, key); // smallest key above the bucket: the successor when the bucket ends at the median
            }
        }
    };
    int64_t done = 0;
    if ((reinterpret_cast<uintptr_t>(img) & 15u) == 0) {
        const int64_t n4 = n >> 2;
        const float4 *src = reinterpret_cast<const float4 *>(img);
        int64_t i = tid;
        // four 16-byte loads in flight per thread: with one, 32 resident warps per SM keep only 16 KB in flight and the pass
        // runs at the latency of L2, not at its bandwidth
        for (; i + 3 * nthreads < n4; i += 4 * nthreads) {
            const float4 v0 = st_ld(src + i), v1 = st_ld(src + i + nthreads), v2 = st_ld(src + i + 2 * nthreads), v3 = st_ld(src + i + 3 * nthreads);
            one(v0.x); one(v0.y); one(v0.z); one(v0.w);
            one(v1.x); one(v1.y); one(v1.z); one(v1.w);
            one(v2.x); one(v2.y); one(v2.z); one(v2.w);
            one(v3.x); one(v3.y); one(v3.z); one(v3.w);
        }
        for (; i < n4; i += nthreads) {
            const float4 v = st_ld(src + i);
            one(v.x); one(v.y); one(v.z); one(v.w);
        }
        done = n4 << 2;
    }
    for (int64_t i = done + tid; i < n; i += nthreads) one(img[i]);
    rc.flush(my);
    if (SHIFT == 8) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) above = min(above, __shfl_xor_sync(0xffffffffu, above, o));
        if ((threadIdx.x & 31) == 0 && above != 0xFFFFFFFFu) atomicMin(&state->above_min, above);
    }
    __syncthreads();
    for (int b = threadIdx.x; b < 256; b += blockDim.x) {
        unsigned t = 0u;
#pragma unroll
        for (int w = 0; w < kHistWarps; ++w) t += s_hist[w][b];
        if (t) atomicAdd(&hist[b], t);
    }
    if (SHIFT == 8) {
        const unsigned staged = min(s_ccount, (unsigned)kStatCandSmem);
        if (threadIdx.x == 0) s_base = staged ? atomicAdd(&state->cand_count, staged) : 0u;
        __syncthreads();
        const unsigned base = s_base;
        for (unsigned t = threadIdx.x; t < staged; t += blockDim.x)
            if (base + t < cand_cap) cand[base + t] = s_cand[t]; // past the cap: cand_count > cand_cap tells the last kernel to rescan the image
    }
    // the last CTA to get here picks the bin
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_last = atomicAdd(&state->ticket, 1u) == gridDim.x - 1u ? 1u : 0u;
    __syncthreads();
    if (s_last && threadIdx.x < 32) {
        __threadfence();
        stat_pick<true>(state, hist, SHIFT, threadIdx.x);
        if (threadIdx.x == 0) state->ticket = 0u;
    }
}

// Last digit, even-count successor and the value, over the candidate buffer (or, when the bucket outgrew the buffer --
// millions of keys sharing 16 bits, e.g. an image that is half zeros -- over the image again; slower, same answer).
__global__ void __launch_bounds__(1024) stat_finish_kernel(const float *__restrict__ img, int64_t n, StatParams p, const StatState *prev,
                                                            StatState *state, const unsigned *__restrict__ cand, unsigned cand_cap) {
    __shared__ unsigned s_hist[256];
    __shared__ unsigned s_min[32];
    if (state->error) {
        if (threadIdx.x == 0) state->result = __int_as_float(0x7fc00000);
        return;
    }
    if (threadIdx.x < 256) s_hist[threadIdx.x] = 0u;
    __syncthreads();
    const unsigned count = state->cand_count;
    const bool rescan = count > cand_cap;
    const int64_t items = rescan ? n : (int64_t)count;
    const float vmed = p.use_vmed ? prev->result : 0.0f;
    const unsigned bucket = state->prefix & 0xFFFF0000u;
    auto key_at = [&](int64_t i, unsigned *key) -> bool {
        if (!rescan) { *key = cand[i]; return true; }
        return stat_value(img[i], p, vmed, key) && (*key & 0xFFFF0000u) == bucket;
    };
    {
        const unsigned prefix = state->prefix, mask = state->mask; // 24 bits
        int64_t i = threadIdx.x;
        if (!rescan) { // one CTA: its speed is the number of loads it keeps in flight
            for (; i + 7 * blockDim.x < items; i += 8 * blockDim.x) {
                unsigned k8[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) k8[u] = __ldg(cand + i + u * blockDim.x);
#pragma unroll
                for (int u = 0; u < 8; ++u)
                    if ((k8[u] & mask) == prefix) atomicAdd(&s_hist[k8[u] & 255u], 1u);
            }
        }
        for (; i < items; i += blockDim.x) {
            unsigned key;
            if (key_at(i, &key) && (key & mask) == prefix) atomicAdd(&s_hist[key & 255u], 1u);
        }
    }
    __syncthreads();
    if (threadIdx.x < 32) stat_pick<false>(state, s_hist, 0, threadIdx.x);
    __syncthreads();
    const unsigned lo = state->prefix; // the key of data[mid - 1] (even count) or data[mid]
    float result = st_unkey(lo);
    if (p.raw) result = stat_normalise(result, p);
    if (state->even) { // (data[mid] + data[mid - 1]) / 2.0: the same key when the run of equal keys continues, else the successor
        unsigned hi = lo;
        if (!(state->kth + 1ull < state->eq_count)) {
            unsigned m = 0xFFFFFFFFu;
            int64_t i = threadIdx.x;
            if (!rescan) {
                for (; i + 7 * blockDim.x < items; i += 8 * blockDim.x) {
                    unsigned k8[8];
#pragma unroll
                    for (int u = 0; u < 8; ++u) k8[u] = __ldg(cand + i + u * blockDim.x);
#pragma unroll
                    for (int u = 0; u < 8; ++u)
                        if (k8[u] > lo && k8[u] < m) m = k8[u];
                }
            }
            for (; i < items; i += blockDim.x) {
                unsigned key;
                if (key_at(i, &key) && key > lo && key < m) m = key;
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) m = min(m, __shfl_xor_sync(0xffffffffu, m, o));
            if ((threadIdx.x & 31) == 0) s_min[threadIdx.x >> 5] = m;
            __syncthreads();
            if (threadIdx.x < 32) {
                m = s_min[threadIdx.x];
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) m = min(m, __shfl_xor_sync(0xffffffffu, m, o));
                hi = min(m, state->above_min); // nothing above it in the bucket: the smallest key of the buckets above
            }
        }
        float a = st_unkey(hi), b = st_unkey(lo);
        if (p.raw) { a = stat_normalise(a, p); b = stat_normalise(b, p); }
        result = __fdiv_rn(__fadd_rn(a, b), 2.0f);
    }
    if (threadIdx.x == 0) state->result = result;
}

// Enqueues the whole selection for one statistic; `state` must be zero-initialised except above_min = 0xFFFFFFFF,
// `hist` 256 zeroed words, `cand` a buffer of cand_cap keys.
cudaError_t stat_select_launch(const float *img, int64_t n, const StatParams &p, const StatState *prev, StatState *state, unsigned *hist,
                               unsigned *cand, unsigned cand_cap, int sm_count, cudaStream_t s) {
    int64_t need = ((n + 3) / 4 + 255) / 256;
    const int64_t cap = (int64_t)sm_count * 4;
    const int grid = (int)(need < 1 ? 1 : (need < cap ? need : cap));
    if (p.raw) {
        stat_pass_kernel<24, true><<<grid, 256, 0, s>>>(img, n, p, prev, state, hist, cand, cand_cap);
        stat_pass_kernel<16, true><<<grid, 256, 0, s>>>(img, n, p, prev, state, hist, cand, cand_cap);
        stat_pass_kernel<8, true><<<grid, 256, 0, s>>>(img, n, p, prev, state, hist, cand, cand_cap);
    } else {
        stat_pass_kernel<24, false><<<grid, 256, 0, s>>>(img, n, p, prev, state, hist, cand, cand_cap);
        stat_pass_kernel<16, false><<<grid, 256, 0, s>>>(img, n, p, prev, state, hist, cand, cand_cap);
        stat_pass_kernel<8, false><<<grid, 256, 0, s>>>(img, n, p, prev, state, hist, cand, cand_cap);
    }
    count_launch();
    count_launch();
    count_launch();
    stat_finish_kernel<<<1, 1024, 0, s>>>(img, n, p, prev, state, cand, cand_cap);
    count_launch();
    return cudaGetLastError();
}

// ---- the same order statistics in ONE pass over the image (round 2) -----------------------------------------------------
// The radix select above reads the image three times per statistic because each pass only learns 8 bits of the answer.  For
// images of a megapixel and more a SAMPLE says where the answer is before the first pass:
//   stat_sample_kernel          one CTA sorts the keys of 4096 jittered stratified pixels (bitonic sort in shared memory, ~3 us)
//                               and brackets the wanted rank: [lo, hi] = the sample keys of rank mid -+ 4 sigma -- the bracket
//                               holds ~6 % of the image and misses the rank about once in 16000 images;
//   stat_bracket_pass_kernel    the ONE pass: counts the kept keys and the keys below lo, histograms the keys of the bracket
//                               over 2048 equal-width bins, copies them to the candidate buffer, tracks the smallest key above
//                               hi; the last CTA to finish (ticket) picks the bin that holds the rank;
//   stat_bracket_finish_kernel  a few CTAs filter the candidates down to that bin (~500 keys); the last one sorts them in
//                               shared memory and writes the value (and the even-count successor).
// Exactness never depends on the sample: a bracket that misses the rank, a candidate buffer or a bin that overflows (heavy
// ties) sets `fallback`, and the host runs the radix select.  2 launches + 1 image pass per statistic instead of 3 passes.
// ---------------------------------------------------------------------------------------------------------------------------
constexpr int kBrSample = 4096;
constexpr int kBrBins = 2048;
constexpr int kBrBinBuf = 8192; // keys of the chosen bin the finish can sort in shared memory

// the key of rank `rank` among the keys the CTA sees through key_at(i), i < count (registers or shared memory): radix select,
// 8 bits per step.  Every thread of the CTA calls it.  s_hist: one 256-bin histogram PER WARP (the keys of an image share
// their leading bits: on one shared histogram the first steps put thousands of atomics on a single address; private to a warp,
// a conflict costs at most the 32 lanes of one instruction), s_pick: 2 words.
// A GROUP of threads (gsize of them, a multiple of 32 and at least 256; gtid = index in the group) runs the select on named
// barrier `bar`, so two groups of one CTA can select two ranks at the same time.  s_hist: min(warps of the group, 8) x 256 words.
template <class KeyAt>
__device__ unsigned stat_group_select(KeyAt key_at, unsigned count, unsigned rank, unsigned *s_hist, unsigned *s_pick, int gtid, int gsize, int bar) {
    unsigned prefix = 0u, mask = 0u;
    const int lane = gtid & 31, nwarp = gsize >> 5, nh = nwarp < 8 ? nwarp : 8, mine = (gtid >> 5) % nh;
    auto sync = [&]() { asm volatile("bar.sync %0, %1;" ::"r"(bar), "r"(gsize) : "memory"); };
    for (int shift = 24; shift >= 0; shift -= 8) {
        for (int i = gtid; i < nh * 256; i += gsize) s_hist[i] = 0u;
        sync();
        for (unsigned i = gtid; i < count; i += gsize) {
            const unsigned key = key_at(i);
            if ((key & mask) == prefix) atomicAdd(&s_hist[mine * 256 + ((key >> shift) & 255u)], 1u);
        }
        sync();
        if (gtid < 256) { // bin gtid summed over the private histograms, left in the first one
            unsigned t = 0u;
            for (int w = 0; w < nh; ++w) t += s_hist[w * 256 + gtid];
            s_hist[gtid] = t;
        }
        sync();
        if (gtid < 32) {
            unsigned c[8], tot = 0u;
#pragma unroll
            for (int j = 0; j < 8; ++j) { c[j] = s_hist[lane * 8 + j]; tot += c[j]; }
            unsigned incl = tot;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned t = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += t;
            }
            unsigned excl = incl - tot;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                if (rank >= excl && rank < excl + c[j]) { s_pick[0] = (unsigned)(lane * 8 + j); s_pick[1] = rank - excl; } // one lane, one j
                excl += c[j];
            }
        }
        sync();
        prefix |= s_pick[0] << shift;
        mask |= 255u << shift;
        rank = s_pick[1];
    }
    return prefix;
}
// the whole CTA as one group (barrier 0 with every thread is __syncthreads)
template <class KeyAt>
__device__ unsigned stat_cta_select(KeyAt key_at, unsigned count, unsigned rank, unsigned *s_hist, unsigned *s_pick) {
    return stat_group_select(key_at, count, rank, s_hist, s_pick, (int)threadIdx.x, (int)blockDim.x, 0);
}

// Limits that stayed on the device (p.mm = the three words minmax_kernel / a reduction's range memo leave: {~key(min), key(max),
// any non-NaN}): decode them and decide the raw-key shortcut here, exactly as the host does for limits it has read back
// (op_min_max, vmed_vmad_run).  The chain then needs no read-back before its last kernel.
__device__ __forceinline__ void stat_resolve(StatParams &p) {
    if (!p.mm) return;
    const unsigned a = __ldcg(p.mm + 0), b = __ldcg(p.mm + 1), any = __ldcg(p.mm + 2);
    const float qnan = __int_as_float(0x7fc00000), inf = __int_as_float(0x7f800000);
    p.vmin = any ? st_unkey(~a) : qnan;
    p.vmax = any ? st_unkey(b) : qnan;
    const float r = __fsub_rn(p.vmax, p.vmin);
    p.raw = (!p.use_vmed && fabsf(p.vmin) < inf && fabsf(p.vmax) < inf && fabsf(r) < inf && p.vmax > p.vmin) ? 1 : 0;
}

__global__ void __launch_bounds__(1024) stat_sample_kernel(const float *__restrict__ img, int64_t n, StatParams p, const StatState *prev,
                                                            StatBracket *br) {
    __shared__ unsigned s_key[kBrSample];
    __shared__ unsigned s_hist[2][8 * 256], s_pick[2][2], s_res[2], s_ms;
    const int tid = threadIdx.x;
    stat_resolve(p);
    const float vmed = p.use_vmed ? prev->result : 0.0f;
    const int64_t stride = n / kBrSample; // n >= kBrSample (host)
    if (tid == 0) { s_ms = 0u; s_res[0] = 0u; s_res[1] = 0xFFFFFFFEu; }
    __syncthreads();
    float v[kBrSample / 1024];
#pragma unroll
    for (int u = 0; u < kBrSample / 1024; ++u) {
        const int j = tid + u * 1024;
        unsigned h = ((unsigned)j * 0x9E3779B9u + 0x7F4A7C15u) * 0x85EBCA6Bu;
        h ^= h >> 15;
        v[u] = img[(int64_t)j * stride + (int64_t)(h % (unsigned)(stride < 0x7FFFFFFF ? stride : 0x7FFFFFFF))];
    }
    unsigned kept = 0u;
#pragma unroll
    for (int u = 0; u < kBrSample / 1024; ++u) {
        unsigned key = 0xFFFFFFFFu, k2; // a dropped sample (NaN, infinite normalised value): ranks last; no kept value has this key
        if (stat_value(v[u], p, vmed, &k2)) { key = k2; ++kept; }
        s_key[tid + u * 1024] = key;
    }
    kept = __reduce_add_sync(0xffffffffu, kept);
    if ((tid & 31) == 0 && kept) atomicAdd(&s_ms, kept);
    __syncthreads();
    const int ms = (int)s_ms;
    if (ms >= 64) { // CTA-uniform
        const int mid = ms / 2;
        const int w = (int)(2.0f * sqrtf((float)ms)) + 2; // 4 sigma of the sample rank of the image's median (sigma = sqrt(ms) / 2)
        // the two halves of the CTA select the two ranks at the same time, each over all the keys, on its own named barrier
        const int grp = tid >> 9, gtid = tid & 511;
        const int rank = grp == 0 ? mid - w : mid + w;
        if (rank >= 0 && rank <= ms - 1) { // uniform within a group
            const unsigned key = stat_group_select([&](unsigned i) { return s_key[i]; }, (unsigned)kBrSample, (unsigned)rank, s_hist[grp], s_pick[grp], gtid, 512, 1 + grp);
            if (gtid == 0) s_res[grp] = key;
        }
    }
    __syncthreads();
    if (tid == 0) {
        const unsigned lo = s_res[0], hi = s_res[1];
        const unsigned span = hi - lo;
        const int bits = 32 - __clz(span | 1u);
        br->lo = lo;
        br->hi = hi;
        br->shift = bits > 11 ? bits - 11 : 0; // (span >> shift) < 2048
    }
}

constexpr int kBrPriv = 16; // candidates a thread stages privately (it meets ~7); more go straight to the buffer

template <bool RAW>
__global__ void __launch_bounds__(256) stat_bracket_pass_kernel(const float *__restrict__ img, int64_t n, StatParams p, const StatState *prev,
                                                                 StatBracket *br, unsigned *hist, unsigned *cand, unsigned cand_cap) {
    __shared__ unsigned s_hist[kBrBins];
    __shared__ unsigned s_priv[kBrPriv][256]; // slot-major: a warp's appends hit 32 banks
    __shared__ unsigned s_last;
    __shared__ unsigned long long s_n, s_below;
    stat_resolve(p);
    // limits decided on the device: both instantiations are launched, the one whose case it is not leaves before touching anything
    // (grid-uniform, ahead of the first barrier: its CTAs retire at once)
    if (p.mm && RAW != (p.raw != 0)) return;
    for (int i = threadIdx.x; i < kBrBins; i += blockDim.x) s_hist[i] = 0u;
    if (threadIdx.x == 0) { s_n = 0ull; s_below = 0ull; }
    __syncthreads();
    const unsigned lo = br->lo, hi = br->hi;
    const int shift = br->shift;
    const bool collect = lo != hi; // one key value in the bracket: its count is all the finish needs
    const float vmed = p.use_vmed ? prev->result : 0.0f;
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nthreads = (int64_t)gridDim.x * blockDim.x;
    unsigned above = 0xFFFFFFFFu, n_seen = 0u, n_drop = 0u, n_below = 0u, pc = 0u;
    const unsigned hist_saddr = (unsigned)__cvta_generic_to_shared(s_hist);
    const unsigned priv_saddr = (unsigned)__cvta_generic_to_shared(&s_priv[0][threadIdx.x]);
    const unsigned span = hi - lo;
    // The common case -- a kept key outside the bracket, 94 % of the pixels -- is kept short: key, one subtraction, three compares,
    // two predicated updates, one branch (14 instructions on raw keys).  Measured with ncu in round 2 (profiles/
    // r2_stats_one_pass_ncu_before.json -> r2_stats_one_pass_ncu.json): 41 -> 34.6 instructions per pixel on raw keys, 63 -> 48 on
    // normalised values; the normalised pass went from 54.2 to 46.6 us, the raw pass stayed at 40 us with 59 % of its issue
    // slots busy -- that one waits on its loads and branches (32 warps per SM), not on issue.  A dropped value (NaN, infinite
    // normalised value) leaves at once and is counted; the kept count is what is left of the pixels seen.
    // Only a bracket key (key - lo <= hi - lo: the subtraction wraps beyond the span for a key below lo) takes the histogram
    // reduction and the private staging column -- no shared atomic with a result anywhere (a returned slot serialises).
    auto one = [&](float v) {
        unsigned key = 0u;
        bool kept;
        if (RAW) {
            kept = !(v != v);
            key = st_key(v);
        } else { // stat_value without its run-time p.raw case: this instantiation only runs on normalised values
            float d = stat_normalise(v, p);
            if (p.use_vmed) d = fabsf(__fsub_rn(d, vmed));
            kept = fabsf(d) < __int_as_float(0x7f800000); // neither NaN nor infinite
            key = st_key(d);
        }
        if (!kept) {
            ++n_drop;
            return;
        }
        const unsigned d = key - lo;
        asm("{\n\t"
            ".reg .pred q;\n\t"
            "setp.lt.u32 q, %1, %2;\n\t"
            "@q add.u32 %0, %0, 1;\n\t"
            "}"
            : "+r"(n_below)
            : "r"(key), "r"(lo));
        if (key > hi) above = min(above, key);
        if (d <= span) {
            asm volatile("red.shared.add.u32 [%0], 1;" ::"r"(hist_saddr + ((d >> shift) << 2)) : "memory");
            if (collect) {
                if (pc < (unsigned)kBrPriv) {
                    asm volatile("st.shared.u32 [%0], %1;" ::"r"(priv_saddr + pc * (256u * 4u)), "r"(key) : "memory");
                } else { // a thread with more than 16 candidates: rare
                    const unsigned at = atomicAdd(&br->cand_count, 1u);
                    if (at < cand_cap) cand[at] = key;
                }
                ++pc;
            }
        }
    };
    int64_t done = 0;
    if ((reinterpret_cast<uintptr_t>(img) & 15u) == 0) {
        const int64_t n4 = n >> 2;
        const float4 *src = reinterpret_cast<const float4 *>(img);
        int64_t i = tid;
        for (; i + 3 * nthreads < n4; i += 4 * nthreads) {
            const float4 v0 = st_ld(src + i), v1 = st_ld(src + i + nthreads), v2 = st_ld(src + i + 2 * nthreads), v3 = st_ld(src + i + 3 * nthreads);
            one(v0.x); one(v0.y); one(v0.z); one(v0.w);
            one(v1.x); one(v1.y); one(v1.z); one(v1.w);
            one(v2.x); one(v2.y); one(v2.z); one(v2.w);
            one(v3.x); one(v3.y); one(v3.z); one(v3.w);
            n_seen += 16u;
        }
        for (; i < n4; i += nthreads) {
            const float4 v = st_ld(src + i);
            one(v.x); one(v.y); one(v.z); one(v.w);
            n_seen += 4u;
        }
        done = n4 << 2;
    }
    for (int64_t i = done + tid; i < n; i += nthreads) {
        one(img[i]);
        n_seen += 1u;
    }
    unsigned n_kept = n_seen - n_drop;
    {
        // this warp's staged candidates -> the buffer: one reservation per warp
        const int lane = threadIdx.x & 31;
        const unsigned mine = min(pc, (unsigned)kBrPriv);
        unsigned incl = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        const unsigned wtot = __shfl_sync(0xffffffffu, incl, 31);
        unsigned base = 0u;
        if (lane == 31 && wtot) base = atomicAdd(&br->cand_count, wtot);
        base = __shfl_sync(0xffffffffu, base, 31) + incl - mine;
        for (unsigned i = 0; i < mine; ++i)
            if (base + i < cand_cap) cand[base + i] = s_priv[i][threadIdx.x];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        above = min(above, __shfl_xor_sync(0xffffffffu, above, o));
        n_kept += __shfl_xor_sync(0xffffffffu, n_kept, o);
        n_below += __shfl_xor_sync(0xffffffffu, n_below, o);
    }
    if ((threadIdx.x & 31) == 0) {
        if (above != 0xFFFFFFFFu) atomicMin(&br->above_min, above);
        atomicAdd(&s_n, (unsigned long long)n_kept);
        atomicAdd(&s_below, (unsigned long long)n_below);
    }
    __syncthreads();
    for (int b = threadIdx.x; b < kBrBins; b += blockDim.x)
        if (s_hist[b]) atomicAdd(&hist[b], s_hist[b]);
    if (threadIdx.x == 0) {
        atomicAdd(&br->n, s_n);
        atomicAdd(&br->below, s_below);
    }
    // the last CTA to get here picks the bin
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_last = atomicAdd(&br->ticket, 1u) == gridDim.x - 1u ? 1u : 0u;
    __syncthreads();
    if (!s_last || threadIdx.x >= 32) return;
    __threadfence();
    const int lane = threadIdx.x;
    const unsigned long long total = *reinterpret_cast<volatile unsigned long long *>(&br->n);
    const unsigned long long below = *reinterpret_cast<volatile unsigned long long *>(&br->below);
    if (total == 0ull) {
        if (lane == 0) br->error = 1; // "Empty data!" (lims.rs:53, 92)
        return;
    }
    const int even = (total % 2ull) == 0ull;
    const unsigned long long kth = even ? total / 2ull - 1ull : total / 2ull;
    constexpr int kPer = kBrBins / 32;
    unsigned hc[kPer]; // all loads in flight at once
#pragma unroll
    for (int j = 0; j < kPer; ++j) hc[j] = __ldcg(&hist[j * 32 + lane]); // bin j * 32 + lane: coalesced
    // lane-major order for the scan: bin b = lane * kPer + j  <->  loaded as hist[j * 32 + lane]; transpose through shared memory
    __shared__ unsigned s_bins[kBrBins];
#pragma unroll
    for (int j = 0; j < kPer; ++j) s_bins[j * 32 + lane] = hc[j];
    __syncwarp();
    unsigned long long tot = 0ull;
    for (int j = 0; j < kPer; ++j) tot += s_bins[lane * kPer + j];
    unsigned long long incl = tot;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned long long t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    const unsigned long long in_count = __shfl_sync(0xffffffffu, incl, 31);
    const unsigned ccount = *reinterpret_cast<volatile unsigned *>(&br->cand_count);
    const bool collected = br->lo != br->hi;
    // the rank must fall inside the bracket, and its candidates must all be in the buffer
    const bool ok = kth >= below && kth < below + in_count && (!collected || (ccount <= cand_cap && (unsigned long long)ccount == in_count));
    if (!ok) {
        if (lane == 0) { br->fallback = 1; br->even = even; }
        return;
    }
    const unsigned long long want = kth - below;
    unsigned long long excl = incl - tot;
    int bin = -1;
    unsigned long long rank_in = 0ull, bcount = 0ull;
    for (int j = 0; j < kPer; ++j) {
        const unsigned long long c = s_bins[lane * kPer + j];
        if (bin < 0 && want < excl + c) { bin = lane * kPer + j; rank_in = want - excl; bcount = c; }
        excl += c;
    }
    const unsigned ballot = __ballot_sync(0xffffffffu, bin >= 0);
    const int src = __ffs(ballot) - 1;
    bin = __shfl_sync(0xffffffffu, bin, src);
    rank_in = __shfl_sync(0xffffffffu, rank_in, src);
    bcount = __shfl_sync(0xffffffffu, bcount, src);
    if (lane == 0) {
        br->bin = (unsigned)bin;
        br->rank_in_bin = rank_in;
        br->bin_count = bcount;
        br->even = even;
        br->in_count = in_count;
        if (collected && bcount > (unsigned long long)(kBrBinBuf - 16 * 256)) br->fallback = 1; // heavy ties in one bin (the finish keeps 16 histograms behind the keys)
    }
}

// grid of a few CTAs over the candidates; the last one to finish sorts the chosen bin's keys and writes the result
__global__ void __launch_bounds__(512) stat_bracket_finish_kernel(StatParams p, StatBracket *br, StatState *state, const unsigned *__restrict__ cand,
                                                                    unsigned *binbuf) {
    __shared__ unsigned s_key[kBrBinBuf];
    __shared__ unsigned s_last;
    const int tid = threadIdx.x;
    stat_resolve(p);
    if (br->error || br->fallback) {
        if (blockIdx.x == 0 && tid == 0) {
            state->error = br->error;
            state->result = __int_as_float(0x7fc00000);
        }
        return;
    }
    const unsigned lo = br->lo, hi = br->hi;
    const int shift = br->shift;
    const unsigned bin = br->bin;
    const int even = br->even;
    unsigned klo, khi;
    if (lo == hi) { // one key value in the bracket
        if (blockIdx.x != 0) return;
        klo = lo;
        khi = (br->rank_in_bin + 1ull < br->bin_count) ? lo : br->above_min;
    } else {
        const unsigned count = br->cand_count;
        unsigned next = 0xFFFFFFFFu;
        const unsigned nthr = gridDim.x * blockDim.x;
        for (unsigned i0 = blockIdx.x * blockDim.x + tid; i0 < count; i0 += 8u * nthr) { // 8 loads in flight per thread
            unsigned k8[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) k8[u] = i0 + u * nthr < count ? __ldg(cand + i0 + u * nthr) : lo + (bin << shift); // filler: in no bin above, dropped below
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                if (!(i0 + u * nthr < count)) continue;
                const unsigned key = k8[u];
                const unsigned b = (key - lo) >> shift;
                if (b == bin) {
                    const unsigned at = atomicAdd(&br->bin_cand_count, 1u);
                    if (at < (unsigned)kBrBinBuf) binbuf[at] = key;
                } else if (b > bin) {
                    next = min(next, key);
                }
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) next = min(next, __shfl_xor_sync(0xffffffffu, next, o));
        if ((tid & 31) == 0 && next != 0xFFFFFFFFu) atomicMin(&br->next_min, next);
        __threadfence();
        __syncthreads();
        if (tid == 0) s_last = atomicAdd(&br->ticket2, 1u) == gridDim.x - 1u ? 1u : 0u;
        __syncthreads();
        if (!s_last) return;
        __threadfence();
        const unsigned c = *reinterpret_cast<volatile unsigned *>(&br->bin_cand_count); // == bin_count <= kBrBinBuf
        const unsigned r = (unsigned)br->rank_in_bin;
        unsigned *const s_hist = s_key + kBrBinBuf - 16 * 256; // 16 warp histograms at the end of the key area (c <= kBrBinBuf - 4096 is checked by the pass)
        __shared__ unsigned s_pick[2], s_cnt_le, s_succ;
        for (unsigned i = tid; i < c; i += blockDim.x) s_key[i] = __ldcg(binbuf + i);
        if (tid == 0) { s_cnt_le = 0u; s_succ = 0xFFFFFFFFu; }
        __syncthreads();
        klo = stat_cta_select([&](unsigned i) { return s_key[i]; }, c, r, s_hist, s_pick);
        khi = klo;
        if (even) { // data[mid]: the same key when the run of equal keys continues, else the smallest key above it
            unsigned le = 0u, succ = 0xFFFFFFFFu;
            for (unsigned i = tid; i < c; i += blockDim.x) {
                const unsigned key = s_key[i];
                le += key <= klo ? 1u : 0u;
                if (key > klo) succ = min(succ, key);
            }
            le = __reduce_add_sync(0xffffffffu, le);
            succ = __reduce_min_sync(0xffffffffu, succ);
            if ((tid & 31) == 0) {
                if (le) atomicAdd(&s_cnt_le, le);
                if (succ != 0xFFFFFFFFu) atomicMin(&s_succ, succ);
            }
            __syncthreads();
            if (!(r + 1u < s_cnt_le)) {
                khi = s_succ;
                if (khi == 0xFFFFFFFFu) khi = min(*reinterpret_cast<volatile unsigned *>(&br->next_min), br->above_min); // the next bin's smallest key, or the smallest key above the bracket
            }
        }
    }
    if (tid == 0) {
        float a = st_unkey(khi), b = st_unkey(klo);
        if (p.raw) { a = stat_normalise(a, p); b = stat_normalise(b, p); }
        state->result = even ? __fdiv_rn(__fadd_rn(a, b), 2.0f) : b;
        state->error = 0;
    }
}

// Enqueues the one-pass selection; `br` zero-initialised except above_min = next_min = 0xFFFFFFFF, `hist` kBrBins zeroed words,
// `cand` cand_cap keys, `binbuf` kBrBinBuf keys.  state->result is valid unless br->fallback (then run stat_select_launch).
cudaError_t stat_bracket_launch(const float *img, int64_t n, const StatParams &p, const StatState *prev, StatState *state, StatBracket *br,
                                unsigned *hist, unsigned *cand, unsigned cand_cap, unsigned *binbuf, int sm_count, cudaStream_t s) {
    if (n < kBrSample) return cudaErrorInvalidValue;
    static const bool dbg = getenv("AFK_STATS_DEBUG") != nullptr; // synchronise after every launch and say which one failed
    auto stage = [&](const char *what) {
        if (!dbg) return;
        const cudaError_t e = cudaStreamSynchronize(s);
        fprintf(stderr, "[afk stats] %s (use_vmed %d): %s\n", what, p.use_vmed, cudaGetErrorString(e));
    };
    stat_sample_kernel<<<1, 1024, 0, s>>>(img, n, p, prev, br);
    count_launch();
    stage("stat_sample_kernel");
    int64_t need = ((n + 3) / 4 + 255) / 256;
    const int64_t cap = (int64_t)sm_count * 4;
    const int grid = (int)(need < 1 ? 1 : (need < cap ? need : cap));
    if (p.mm && !p.use_vmed) { // raw keys or normalised values is decided on the device (stat_resolve): one of the two returns at once
        stat_bracket_pass_kernel<true><<<grid, 256, 0, s>>>(img, n, p, prev, br, hist, cand, cand_cap);
        stat_bracket_pass_kernel<false><<<grid, 256, 0, s>>>(img, n, p, prev, br, hist, cand, cand_cap);
        count_launch();
    } else if (p.raw && !p.mm) {
        stat_bracket_pass_kernel<true><<<grid, 256, 0, s>>>(img, n, p, prev, br, hist, cand, cand_cap);
    } else {
        stat_bracket_pass_kernel<false><<<grid, 256, 0, s>>>(img, n, p, prev, br, hist, cand, cand_cap);
    }
    count_launch();
    stage("stat_bracket_pass_kernel");
    // (One kernel that branches on p.raw between the two instantiations of the stream was tried in round 2 and dropped: inlined
    // side by side, the PTX nvcc 12.9 emits for the non-RAW copy lacks the `key > hi` term of the bracket test, and the
    // histogram reduction then lands outside shared memory.  Two launches, one of which returns at once, cost ~6 us.)
    stat_bracket_finish_kernel<<<128, 512, 0, s>>>(p, br, state, cand, binbuf);
    count_launch();
    stage("stat_bracket_finish_kernel");
    return cudaGetLastError();
}

// ---- zscale sample (zscale.rs:101-126): every stride-th pixel of every stride-th row of the flipped image ---------------
__global__ void zsample_kernel(const float *__restrict__ img, int64_t ny, int64_t nx, int64_t stride, int64_t rows_s, int64_t cols_s,
                               float *out) {
    const int64_t total = rows_s * cols_s;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / cols_s, c = i - r * cols_s;
        const int64_t y = ny - 1 - r * stride; // coerce_to_array_view2 flips the rows (image.rs:123-129)
        out[i] = img[y * nx + c * stride];
    }
}

cudaError_t zsample_launch(const float *img, int64_t ny, int64_t nx, int64_t stride, int64_t rows_s, int64_t cols_s, float *out,
                           cudaStream_t s) {
    const int64_t total = rows_s * cols_s;
    if (total <= 0) return cudaSuccess;
    zsample_kernel<<<(unsigned)((total + 255) / 256), 256, 0, s>>>(img, ny, nx, stride, rows_s, cols_s, out);
    count_launch();
    return cudaGetLastError();
}

} // namespace afk
