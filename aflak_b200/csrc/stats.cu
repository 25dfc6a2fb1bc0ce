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

constexpr int kHistWarps = 8;

// counts: [nbins][channels] (bin-major like Vec<[Bin; 3]>), 64-bit
__global__ void __launch_bounds__(256) histogram_kernel(const float *__restrict__ img, int64_t n_per_channel, int channels, float mn,
                                                         float mx, unsigned nb, unsigned long long *counts) {
    extern __shared__ unsigned s_hist[]; // [kHistWarps][nb]
    for (unsigned i = threadIdx.x; i < kHistWarps * nb; i += blockDim.x) s_hist[i] = 0u;
    __syncthreads();
    unsigned *my = s_hist + (threadIdx.x >> 5) * nb;
    const float range = __fsub_rn(mx, mn), nb_f = (float)nb;
    const int c = blockIdx.y;
    const float *src = img + (size_t)c * n_per_channel;
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nthreads = (int64_t)gridDim.x * blockDim.x;
    RunCounter rc;
    int64_t done = 0;
    if ((reinterpret_cast<uintptr_t>(src) & 15u) == 0) {
        const int64_t n4 = n_per_channel >> 2;
        for (int64_t i = tid; i < n4; i += nthreads) {
            const float4 v = st_ld(reinterpret_cast<const float4 *>(src) + i);
            rc.add(hist_bin(v.x, mn, range, nb_f, nb), my);
            rc.add(hist_bin(v.y, mn, range, nb_f, nb), my);
            rc.add(hist_bin(v.z, mn, range, nb_f, nb), my);
            rc.add(hist_bin(v.w, mn, range, nb_f, nb), my);
        }
        done = n4 << 2;
    }
    for (int64_t i = done + tid; i < n_per_channel; i += nthreads) rc.add(hist_bin(src[i], mn, range, nb_f, nb), my);
    rc.flush(my);
    __syncthreads();
    for (unsigned b = threadIdx.x; b < nb; b += blockDim.x) {
        unsigned t = 0u;
#pragma unroll
        for (int w = 0; w < kHistWarps; ++w) t += s_hist[w * nb + b];
        if (t) atomicAdd(&counts[(size_t)b * channels + c], (unsigned long long)t);
    }
}

cudaError_t histogram_launch(const float *img, int64_t n_per_channel, int channels, float mn, float mx, int nbins,
                             unsigned long long *counts_dev, int sm_count, cudaStream_t s) {
    if (n_per_channel <= 0) return cudaSuccess;
    if (nbins < 1 || nbins > 1024 || channels < 1 || channels > 3) return cudaErrorInvalidValue;
    int64_t need = ((n_per_channel + 3) / 4 + 255) / 256;
    const int64_t cap = (int64_t)sm_count * 4;
    const dim3 grid((unsigned)(need < 1 ? 1 : (need < cap ? need : cap)), (unsigned)channels);
    histogram_kernel<<<grid, 256, (size_t)kHistWarps * nbins * sizeof(unsigned), s>>>(img, n_per_channel, channels, mn, mx, (unsigned)nbins, counts_dev);
    count_launch();
    return cudaGetLastError();
}

// ---- lims::get_vmed_normalized / get_vmad_normalized: exact order statistics by radix select ---------------------------
__device__ __forceinline__ unsigned st_key(float f) {
    const unsigned b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float st_unkey(unsigned k) {
    const unsigned b = (k & 0x80000000u) ? (k & 0x7FFFFFFFu) : ~k;
    return __uint_as_float(b);
}

// the value the reference pushes into `data` before sorting: d = (v - vmin) / (vmax - vmin) (lims.rs:48), or
// |d - vmed| (lims.rs:87); kept only when neither NaN nor infinite
__device__ __forceinline__ bool stat_value(float v, const StatParams &p, float vmed, unsigned *key) {
    float d = __fdiv_rn(__fsub_rn(v, p.vmin), __fsub_rn(p.vmax, p.vmin));
    if (p.use_vmed) d = fabsf(__fsub_rn(d, vmed));
    if (!(fabsf(d) < __int_as_float(0x7f800000))) return false; // NaN or +-Inf
    *key = st_key(d);
    return true;
}

__global__ void __launch_bounds__(256) stat_hist_kernel(const float *__restrict__ img, int64_t n, StatParams p, const StatState *prev,
                                                         const StatState *state, unsigned *hist, int shift) {
    __shared__ unsigned s_hist[kHistWarps][256];
    for (int i = threadIdx.x; i < kHistWarps * 256; i += blockDim.x) (&s_hist[0][0])[i] = 0u;
    __syncthreads();
    if (state->error) return;
    unsigned *my = s_hist[threadIdx.x >> 5];
    const unsigned prefix = state->prefix, mask = state->mask;
    const float vmed = p.use_vmed ? prev->result : 0.0f;
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nthreads = (int64_t)gridDim.x * blockDim.x;
    RunCounter rc;
    auto one = [&](float v) {
        unsigned key;
        if (stat_value(v, p, vmed, &key) && (key & mask) == prefix) rc.add((key >> shift) & 255u, my);
    };
    int64_t done = 0;
    if ((reinterpret_cast<uintptr_t>(img) & 15u) == 0) {
        const int64_t n4 = n >> 2;
        for (int64_t i = tid; i < n4; i += nthreads) {
            const float4 v = st_ld(reinterpret_cast<const float4 *>(img) + i);
            one(v.x); one(v.y); one(v.z); one(v.w);
        }
        done = n4 << 2;
    }
    for (int64_t i = done + tid; i < n; i += nthreads) one(img[i]);
    rc.flush(my);
    __syncthreads();
    for (int b = threadIdx.x; b < 256; b += blockDim.x) {
        unsigned t = 0u;
#pragma unroll
        for (int w = 0; w < kHistWarps; ++w) t += s_hist[w][b];
        if (t) atomicAdd(&hist[b], t);
    }
}

// one warp: find the bin holding rank kth among the candidates, narrow the prefix, clear the histogram for the next pass
__global__ void stat_pick_kernel(StatState *state, unsigned *hist, int shift) {
    const int lane = threadIdx.x;
    if (state->error) return;
    unsigned long long c[8], tot = 0ull;
#pragma unroll
    for (int j = 0; j < 8; ++j) { c[j] = hist[lane * 8 + j]; tot += c[j]; }
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
            if (lane == 0) state->error = 1; // "Empty data!" (lims.rs:53, 92)
            return;
        }
        const int even = (total % 2ull) == 0ull;
        kth = even ? total / 2ull - 1ull : total / 2ull; // data[mid - 1] (then data[mid] via `next`) or data[mid]
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
    if (lane == 0) {
        state->kth = kth - below;
        state->eq_count = cnt;
        state->prefix |= (unsigned)digit << shift;
        state->mask |= 255u << shift;
        state->next_key = 0xFFFFFFFFu;
    }
}

// even count: the upper middle element is the same key when the equal run continues, else the smallest key above it
__global__ void __launch_bounds__(256) stat_next_kernel(const float *__restrict__ img, int64_t n, StatParams p, const StatState *prev,
                                                         StatState *state) {
    if (state->error || !state->even || state->kth + 1ull < state->eq_count) return;
    const unsigned lo = state->prefix;
    const float vmed = p.use_vmed ? prev->result : 0.0f;
    unsigned m = 0xFFFFFFFFu;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        unsigned key;
        if (stat_value(img[i], p, vmed, &key) && key > lo && key < m) m = key;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = min(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0 && m != 0xFFFFFFFFu) atomicMin(&state->next_key, m);
}

__global__ void stat_final_kernel(StatState *state) {
    if (state->error) {
        state->result = __int_as_float(0x7fc00000);
        return;
    }
    const float lo = st_unkey(state->prefix);
    if (!state->even) { state->result = lo; return; }
    const float hi = (state->kth + 1ull < state->eq_count) ? lo : st_unkey(state->next_key);
    state->result = __fdiv_rn(__fadd_rn(hi, lo), 2.0f); // (data[mid] + data[mid - 1]) / 2.0
}

// Enqueues the whole selection for one statistic; `state` must be zero-initialised, `hist` 256 zeroed words.
cudaError_t stat_select_launch(const float *img, int64_t n, const StatParams &p, const StatState *prev, StatState *state, unsigned *hist,
                               int sm_count, cudaStream_t s) {
    int64_t need = ((n + 3) / 4 + 255) / 256;
    const int64_t cap = (int64_t)sm_count * 4;
    const int grid = (int)(need < 1 ? 1 : (need < cap ? need : cap));
    for (int shift = 24; shift >= 0; shift -= 8) {
        stat_hist_kernel<<<grid, 256, 0, s>>>(img, n, p, prev, state, hist, shift);
        count_launch();
        stat_pick_kernel<<<1, 32, 0, s>>>(state, hist, shift);
        count_launch();
    }
    stat_next_kernel<<<grid, 256, 0, s>>>(img, n, p, prev, state);
    count_launch();
    stat_final_kernel<<<1, 1, 0, s>>>(state);
    count_launch();
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
