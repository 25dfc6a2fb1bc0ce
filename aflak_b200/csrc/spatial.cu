// spatial.cu -- the operators that reduce or gather over the two spatial axes: extract_wave (sum of a
// region of interest per frame -> 1-D spectrum, the orthogonal reduction to `integral`), extrude, slice_3d_to_2d.
//
// Reference: aflak_primitives/src/lib.rs:1147-1169 (run_extract_wave), :1233-1256 (run_extrude), :928-1128
// (run_slice_3d_to_2d), roi.rs:12-37 (ROI::filter).
#include "kernels.h"

#include <cstdint>

namespace afk {

__device__ __forceinline__ float4 sp_ld(const float4 *p) {
    float4 r;
    asm("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}

// ---------------------------------------------------------------------------------------------------------------
// extract_wave, default mode: out[k] = sum over the region of cube[k][pixel], accumulated in f64.
// grid = (chunks, frames): a CTA sums one chunk of the region for one frame and stores an f64 partial; a second tiny
// kernel adds the partials of a frame in chunk order (deterministic, no float atomics).
// `offsets` == nullptr: the region is the whole frame (ROI::All), read with 16-byte lanes.
// ---------------------------------------------------------------------------------------------------------------
constexpr int kWaveBlock = 256;

__device__ __forceinline__ double block_sum(double v) {
    __shared__ double s[kWaveBlock / 32];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (l == 0) s[w] = v;
    __syncthreads();
    double t = 0.0;
    if (w == 0) {
        t = l < kWaveBlock / 32 ? s[l] : 0.0;
#pragma unroll
        for (int o = 4; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    }
    return t; // valid in thread 0
}

__global__ void __launch_bounds__(kWaveBlock) wave_partial_kernel(const float *__restrict__ cube, int64_t n2, const int64_t *__restrict__ offsets,
                                                                   int64_t count, int64_t chunk, double *partials) {
    const int64_t k = blockIdx.y;
    const int64_t lo = (int64_t)blockIdx.x * chunk, hi = min(count, lo + chunk);
    const float *frame = cube + k * n2;
    double acc = 0.0;
    if (offsets) {
        for (int64_t i = lo + threadIdx.x; i < hi; i += kWaveBlock) acc += (double)frame[offsets[i]];
    } else {
        const bool vec = ((reinterpret_cast<uintptr_t>(frame + lo) & 15u) == 0);
        int64_t i = lo;
        if (vec) {
            const int64_t n4 = (hi - lo) >> 2;
            const float4 *f4 = reinterpret_cast<const float4 *>(frame + lo);
            // four independent 16-byte loads per thread per trip
            int64_t j = threadIdx.x;
            for (; j + 3 * kWaveBlock < n4; j += 4 * kWaveBlock) {
                const float4 a = sp_ld(f4 + j), b = sp_ld(f4 + j + kWaveBlock), c = sp_ld(f4 + j + 2 * kWaveBlock), d = sp_ld(f4 + j + 3 * kWaveBlock);
                acc += ((double)a.x + (double)a.y) + ((double)a.z + (double)a.w);
                acc += ((double)b.x + (double)b.y) + ((double)b.z + (double)b.w);
                acc += ((double)c.x + (double)c.y) + ((double)c.z + (double)c.w);
                acc += ((double)d.x + (double)d.y) + ((double)d.z + (double)d.w);
            }
            for (; j < n4; j += kWaveBlock) {
                const float4 a = sp_ld(f4 + j);
                acc += ((double)a.x + (double)a.y) + ((double)a.z + (double)a.w);
            }
            i = lo + (n4 << 2);
        }
        for (int64_t t = i + threadIdx.x; t < hi; t += kWaveBlock) acc += (double)frame[t];
    }
    const double total = block_sum(acc);
    if (threadIdx.x == 0) partials[k * gridDim.x + blockIdx.x] = total;
}

__global__ void wave_finish_kernel(const double *partials, int chunks, int64_t nz, float *out) {
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nz) return;
    double t = 0.0;
    for (int c = 0; c < chunks; ++c) t += partials[k * chunks + c];
    out[k] = (float)t;
}

// sequential mode: the reference's own order -- res = 0.0; res += val over the region in list order (roi.rs:17-34),
// f32.  One thread per frame; the chain of dependent additions is the floor (region size x 4 cycles).
__global__ void __launch_bounds__(32) wave_sequential_kernel(const float *__restrict__ cube, int64_t n2, int64_t nz,
                                                              const int64_t *__restrict__ offsets, int64_t count, float *out) {
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nz) return;
    const float *frame = cube + k * n2;
    float res = 0.0f;
    if (offsets) {
        int64_t i = 0;
        for (; i + 8 <= count; i += 8) {
            float v[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = frame[offsets[i + j]];
#pragma unroll
            for (int j = 0; j < 8; ++j) res = __fadd_rn(res, v[j]);
        }
        for (; i < count; ++i) res = __fadd_rn(res, frame[offsets[i]]);
    } else {
        int64_t i = 0;
        if ((reinterpret_cast<uintptr_t>(frame) & 15u) == 0) {
            const float4 *f4 = reinterpret_cast<const float4 *>(frame);
            const int64_t n4 = count >> 2;
            int64_t j = 0;
            for (; j + 8 <= n4; j += 8) { // 128 bytes = one L1 line per trip, loaded before the dependent adds
                float4 v[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) v[u] = f4[j + u];
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    res = __fadd_rn(res, v[u].x); res = __fadd_rn(res, v[u].y);
                    res = __fadd_rn(res, v[u].z); res = __fadd_rn(res, v[u].w);
                }
            }
            for (; j < n4; ++j) {
                const float4 v = f4[j];
                res = __fadd_rn(res, v.x); res = __fadd_rn(res, v.y); res = __fadd_rn(res, v.z); res = __fadd_rn(res, v.w);
            }
            i = n4 << 2;
        }
        for (; i < count; ++i) res = __fadd_rn(res, frame[i]);
    }
    out[k] = res;
}

// Sequential mode over whole frames (ROI::All), the common case: the chain of dependent f32 additions per frame is the
// floor (n2 x 4 cycles; 2.2 ms for a 1024 x 1024 frame, the same as the HBM time of the whole cube), so the job is to
// keep it fed.  One warp owns 32 frames, lane = frame.  Reading a frame lane by lane asks HBM for one 16-byte piece per
// lane per instruction (32 different lines) and is bound by the request rate at ~0.75 TB/s; instead the warp copies
// 512 contiguous bytes of ONE frame per cp.async instruction (lane = 16-byte unit), 32 frames per stage, into a ring in
// shared memory, rotated by the frame number so that both the copy (frame fixed, units across lanes) and the read-back
// (unit fixed, frames across lanes) are bank-conflict free.  Each lane of the consumer warp then adds its own frame's
// values in order; a second warp issues the copies (one barrier per stage), which would otherwise take issue slots
// from the chain (4.0 ms with one warp doing both).
constexpr int kWsStages = 6;                  // 6 x 16 KiB in flight per SM: ~3x the bandwidth-latency product of one SM's share
constexpr int kWsStageUnits = 32 * 32;        // float4 units per stage: 32 frames x 512 bytes
__global__ void __launch_bounds__(64) wave_sequential_tiled_kernel(const float *__restrict__ cube, int64_t n2, int64_t nz, float *out) {
    extern __shared__ float4 ws_ring[];
    const int lane = threadIdx.x & 31;
    const bool producer = threadIdx.x >= 32; // warp 1 only issues copies; warp 0 only adds: its issue slots belong to the chain
    const int64_t k0 = (int64_t)blockIdx.x * 32;
    const int64_t nseg = n2 >> 7; // 128 values of every frame per stage
    if (producer) {
        const unsigned ring = (unsigned)__cvta_generic_to_shared(ws_ring);
        auto issue = [&](int64_t seg, int stage) {
            if (seg < nseg) {
                const float *src0 = cube + seg * 128 + lane * 4;
                const unsigned dst0 = ring + (unsigned)(stage * kWsStageUnits) * 16u;
#pragma unroll
                for (int f = 0; f < 32; ++f) {
                    const int64_t fk = k0 + f < nz ? k0 + f : nz - 1; // frames past the end re-read the last one (result dropped)
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst0 + (unsigned)(f * 32 + ((lane + f) & 31)) * 16u),
                                 "l"(src0 + fk * n2)
                                 : "memory");
                }
            }
            asm volatile("cp.async.commit_group;" ::: "memory"); // one group per stage, empty past the end: the wait count stays fixed
        };
#pragma unroll
        for (int st = 0; st < kWsStages - 1; ++st) issue(st, st);
        int stage = kWsStages - 1; // the stage to refill after the barrier of trip `seg`: the one consumed in trip seg - 1
        for (int64_t seg = 0; seg < nseg; ++seg) {
            asm volatile("cp.async.wait_group %0;" ::"n"(kWsStages - 2) : "memory"); // stage of `seg` has landed
            __syncthreads();                                                           // ... and the consumer is done with seg - 1
            issue(seg + kWsStages - 1, stage);
            if (++stage == kWsStages) stage = 0;
        }
        return;
    }
    float res = 0.0f;
    int stage = 0;
    for (int64_t seg = 0; seg < nseg; ++seg) {
        __syncthreads();
        const float4 *src = ws_ring + stage * kWsStageUnits + lane * 32;
#pragma unroll
        for (int c = 0; c < 32; ++c) {
            const float4 v = src[(c + lane) & 31];
            res = __fadd_rn(res, v.x); res = __fadd_rn(res, v.y); res = __fadd_rn(res, v.z); res = __fadd_rn(res, v.w);
        }
        if (++stage == kWsStages) stage = 0;
    }
    const int64_t k = k0 + lane;
    if (k >= nz) return;
    const float *frame = cube + k * n2;
    for (int64_t i = nseg << 7; i < n2; ++i) res = __fadd_rn(res, frame[i]);
    out[k] = res;
}

size_t wave_partial_bytes(int64_t nz, int64_t count, int *chunks_out, int64_t *chunk_out) {
    // 64 Ki pixels per chunk: a 256-thread CTA streams 256 KiB; small regions are a single chunk
    int64_t chunk = 65536;
    int64_t chunks = (count + chunk - 1) / chunk;
    if (chunks < 1) chunks = 1;
    if (chunks > 65535) { // grid.x limit is generous, keep the partial buffer modest anyway
        chunks = 65535;
        chunk = ((count + chunks - 1) / chunks + 3) & ~(int64_t)3;
        chunks = (count + chunk - 1) / chunk;
    }
    *chunks_out = (int)chunks;
    *chunk_out = chunk;
    return (size_t)nz * (size_t)chunks * sizeof(double);
}

cudaError_t wave_launch(const float *cube, int64_t nz, int64_t n2, const int64_t *offsets_dev, int64_t count, int mode, void *partials,
                        float *out, cudaStream_t s) {
    if (nz <= 0) return cudaSuccess;
    if (nz > 65535 * 1024) return cudaErrorInvalidValue;
    if (mode == SUM_SEQUENTIAL) {
        // whole frames, 16-byte aligned, and deep enough that the ring pays: coalesced staging; otherwise lane = frame reads
        if (!offsets_dev && count == n2 && (n2 & 3) == 0 && n2 >= 1024 && (reinterpret_cast<uintptr_t>(cube) & 15u) == 0) {
            constexpr int smem_ws = kWsStages * kWsStageUnits * 16;
            static bool attr_ws[16] = {};
            int dev = 0;
            cudaGetDevice(&dev);
            if (dev >= 0 && dev < 16 && !attr_ws[dev]) {
                const cudaError_t ea = cudaFuncSetAttribute(wave_sequential_tiled_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_ws);
                if (ea != cudaSuccess) return ea;
                attr_ws[dev] = true;
            }
            wave_sequential_tiled_kernel<<<(unsigned)((nz + 31) / 32), 64, smem_ws, s>>>(cube, n2, nz, out);
            count_launch();
            return cudaGetLastError();
        }
        wave_sequential_kernel<<<(unsigned)((nz + 31) / 32), 32, 0, s>>>(cube, n2, nz, offsets_dev, count, out);
        count_launch();
        return cudaGetLastError();
    }
    int chunks;
    int64_t chunk;
    wave_partial_bytes(nz, count, &chunks, &chunk);
    if (!partials) return cudaErrorInvalidValue;
    // frames go in grid.y (<= 65535): loop over slabs of frames for deeper cubes
    for (int64_t k0 = 0; k0 < nz; k0 += 65535) {
        const int64_t kn = nz - k0 < 65535 ? nz - k0 : 65535;
        wave_partial_kernel<<<dim3((unsigned)chunks, (unsigned)kn), kWaveBlock, 0, s>>>(cube + k0 * n2, n2, offsets_dev, count, chunk,
                                                                                        static_cast<double *>(partials) + k0 * chunks);
        count_launch();
    }
    wave_finish_kernel<<<(unsigned)((nz + 255) / 256), 256, 0, s>>>(static_cast<const double *>(partials), chunks, nz, out);
    count_launch();
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------------
// extrude (lib.rs:1233-1256): out[r][p] = cube[nz-1-r][pixel p]  (frames in reverse order, region in list order)
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) extrude_kernel(const float *__restrict__ cube, int64_t nz, int64_t n2, const int64_t *__restrict__ offsets,
                                                       int64_t count, float *__restrict__ out) {
    const int64_t total = nz * count;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / count, p = i - r * count;
        out[i] = cube[(nz - 1 - r) * n2 + offsets[p]];
    }
}

cudaError_t extrude_launch(const float *cube, int64_t nz, int64_t n2, const int64_t *offsets_dev, int64_t count, float *out, int sm_count,
                           cudaStream_t s) {
    const int64_t total = nz * count;
    if (total <= 0) return cudaSuccess;
    int64_t blocks = (total + 255) / 256;
    if (blocks > (int64_t)sm_count * 16) blocks = (int64_t)sm_count * 16;
    extrude_kernel<<<(unsigned)blocks, 256, 0, s>>>(cube, nz, n2, offsets_dev, count, out);
    count_launch();
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------------
// slice_3d_to_2d (lib.rs:1096-1110): out[i] = image[x as usize, y as usize, z as usize] for map[i] = (x, y, z);
// an index outside the cube is an error naming the first such point in map order -> *first_bad = min linear index
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ bool as_index(float f, int64_t dim, int64_t *idx) { // Rust `as usize`: truncate, saturate, NaN -> 0
    int64_t i;
    if (!(f > 0.0f)) i = 0;
    else if (f >= 9.2233720368547758e18f) return false;
    else i = (int64_t)f;
    *idx = i;
    return i < dim;
}

__global__ void __launch_bounds__(256) slice3d_kernel(const float *__restrict__ cube, int64_t d0, int64_t d1, int64_t d2,
                                                       const float *__restrict__ map, int64_t n, float *__restrict__ out,
                                                       unsigned long long *first_bad) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        int64_t a, b, c;
        const bool good = as_index(map[3 * i + 0], d0, &a) & as_index(map[3 * i + 1], d1, &b) & as_index(map[3 * i + 2], d2, &c);
        if (good) out[i] = cube[(a * d1 + b) * d2 + c];
        else {
            out[i] = 0.0f;
            atomicMin(first_bad, (unsigned long long)i);
        }
    }
}

cudaError_t slice3d_launch(const float *cube, int64_t d0, int64_t d1, int64_t d2, const float *map_dev, int64_t n, float *out,
                           unsigned long long *first_bad, int sm_count, cudaStream_t s) {
    if (n <= 0) return cudaSuccess;
    int64_t blocks = (n + 255) / 256;
    if (blocks > (int64_t)sm_count * 16) blocks = (int64_t)sm_count * 16;
    slice3d_kernel<<<(unsigned)blocks, 256, 0, s>>>(cube, d0, d1, d2, map_dev, n, out, first_bad);
    count_launch();
    return cudaGetLastError();
}

} // namespace afk
