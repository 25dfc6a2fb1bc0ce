// ingest.cu -- FITS data unit -> f32 cube on the device.
//
// Replaces the host passes the reference makes after fitrs has read the file (aflak_primitives/src/unit.rs:283-331,
// fits.rs:42-103): big-endian decode, i16/i32/u32/f64 -> f32 conversion, BLANK -> NaN, then `image * bscale + bzero`
// (two separately rounded f32 steps, always applied, as unit.rs:328 does).  The raw big-endian bytes are what crosses
// PCIe; this kernel runs on each uploaded chunk while the next chunk is in flight (runtime: fits.cpp).
//
// HBM-bound: (element size + 4) bytes per voxel, 16-byte lanes, grid-stride.  4-byte formats may be decoded in place
// (out == raw): a thread reads its 16 bytes before it writes them.
#include "kernels.h"

#include <cstdint>

namespace afk {

__device__ __forceinline__ uint32_t bswap32(uint32_t x) { return __byte_perm(x, 0u, 0x0123); }

template <int FMT> struct Ing;
template <> struct Ing<ING_F32> {
    static constexpr int kBytes = 4;
    static __device__ __forceinline__ float one(const uint8_t *p, int64_t i, const IngestArgs &) {
        return __uint_as_float(bswap32(reinterpret_cast<const uint32_t *>(p)[i]));
    }
    static __device__ __forceinline__ void four(const uint8_t *p, int64_t i4, const IngestArgs &, float (&v)[4]) {
        const uint4 r = reinterpret_cast<const uint4 *>(p)[i4];
        v[0] = __uint_as_float(bswap32(r.x)); v[1] = __uint_as_float(bswap32(r.y));
        v[2] = __uint_as_float(bswap32(r.z)); v[3] = __uint_as_float(bswap32(r.w));
    }
};
template <> struct Ing<ING_I32> { // FitsData::IntegersI32(Option<i32>): BLANK -> None -> NaN, else `int as f32` (fits.rs:64-82)
    static constexpr int kBytes = 4;
    static __device__ __forceinline__ float conv(uint32_t be, const IngestArgs &a) {
        const int32_t x = (int32_t)bswap32(be);
        return (a.has_blank && (int64_t)x == a.blank) ? __int_as_float(0x7fc00000) : __int2float_rn(x);
    }
    static __device__ __forceinline__ float one(const uint8_t *p, int64_t i, const IngestArgs &a) {
        return conv(reinterpret_cast<const uint32_t *>(p)[i], a);
    }
    static __device__ __forceinline__ void four(const uint8_t *p, int64_t i4, const IngestArgs &a, float (&v)[4]) {
        const uint4 r = reinterpret_cast<const uint4 *>(p)[i4];
        v[0] = conv(r.x, a); v[1] = conv(r.y, a); v[2] = conv(r.z, a); v[3] = conv(r.w, a);
    }
};
template <> struct Ing<ING_U32> { // FitsData::IntegersU32(Option<u32>) (fits.rs:84-102)
    static constexpr int kBytes = 4;
    static __device__ __forceinline__ float conv(uint32_t be, const IngestArgs &a) {
        const uint32_t x = bswap32(be);
        return (a.has_blank && (int64_t)x == a.blank) ? __int_as_float(0x7fc00000) : __uint2float_rn(x);
    }
    static __device__ __forceinline__ float one(const uint8_t *p, int64_t i, const IngestArgs &a) {
        return conv(reinterpret_cast<const uint32_t *>(p)[i], a);
    }
    static __device__ __forceinline__ void four(const uint8_t *p, int64_t i4, const IngestArgs &a, float (&v)[4]) {
        const uint4 r = reinterpret_cast<const uint4 *>(p)[i4];
        v[0] = conv(r.x, a); v[1] = conv(r.y, a); v[2] = conv(r.z, a); v[3] = conv(r.w, a);
    }
};
template <> struct Ing<ING_I16> { // BITPIX 16: widened to i32 by the reader, then as above
    static constexpr int kBytes = 2;
    static __device__ __forceinline__ float conv(uint32_t be16, const IngestArgs &a) {
        const int32_t x = (int32_t)(int16_t)(uint16_t)(((be16 & 0xFFu) << 8) | ((be16 >> 8) & 0xFFu));
        return (a.has_blank && (int64_t)x == a.blank) ? __int_as_float(0x7fc00000) : __int2float_rn(x);
    }
    static __device__ __forceinline__ float one(const uint8_t *p, int64_t i, const IngestArgs &a) {
        return conv(reinterpret_cast<const uint16_t *>(p)[i], a);
    }
    static __device__ __forceinline__ void four(const uint8_t *p, int64_t i4, const IngestArgs &a, float (&v)[4]) {
        const uint2 r = reinterpret_cast<const uint2 *>(p)[i4];
        v[0] = conv(r.x & 0xFFFFu, a); v[1] = conv(r.x >> 16, a);
        v[2] = conv(r.y & 0xFFFFu, a); v[3] = conv(r.y >> 16, a);
    }
};
template <> struct Ing<ING_F64> { // FitsData::FloatingPoint64: `f as f32` (fits.rs:53-62), round to nearest even
    static constexpr int kBytes = 8;
    static __device__ __forceinline__ float conv(uint32_t w0, uint32_t w1) { // w0 = first 4 bytes of the big-endian double
        const uint64_t bits = ((uint64_t)bswap32(w0) << 32) | (uint64_t)bswap32(w1);
        return __double2float_rn(__longlong_as_double((long long)bits));
    }
    static __device__ __forceinline__ float one(const uint8_t *p, int64_t i, const IngestArgs &) {
        const uint2 r = reinterpret_cast<const uint2 *>(p)[i];
        return conv(r.x, r.y);
    }
    static __device__ __forceinline__ void four(const uint8_t *p, int64_t i4, const IngestArgs &, float (&v)[4]) {
        const uint4 r0 = reinterpret_cast<const uint4 *>(p)[2 * i4], r1 = reinterpret_cast<const uint4 *>(p)[2 * i4 + 1];
        v[0] = conv(r0.x, r0.y); v[1] = conv(r0.z, r0.w); v[2] = conv(r1.x, r1.y); v[3] = conv(r1.z, r1.w);
    }
};

// `image * bscale + bzero` (unit.rs:328): ndarray scalar ops, one rounding each, no FMA
__device__ __forceinline__ float ing_scale(float v, const IngestArgs &a) { return __fadd_rn(__fmul_rn(v, a.bscale), a.bzero); }

template <int FMT> __global__ void __launch_bounds__(256) ingest_kernel(const IngestArgs a, int vec_ok) {
    const uint8_t *raw = static_cast<const uint8_t *>(a.raw);
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t nthreads = (int64_t)gridDim.x * blockDim.x;
    int64_t done = 0;
    if (vec_ok) {
        const int64_t n4 = a.n >> 2;
        float4 *o4 = reinterpret_cast<float4 *>(a.out);
        for (int64_t i = tid; i < n4; i += nthreads) {
            float v[4];
            Ing<FMT>::four(raw, i, a, v);
            o4[i] = make_float4(ing_scale(v[0], a), ing_scale(v[1], a), ing_scale(v[2], a), ing_scale(v[3], a));
        }
        done = n4 << 2;
    }
    for (int64_t i = done + tid; i < a.n; i += nthreads) a.out[i] = ing_scale(Ing<FMT>::one(raw, i, a), a);
}

int ingest_element_bytes(int fmt) {
    switch (fmt) {
    case ING_F32: case ING_I32: case ING_U32: return 4;
    case ING_I16: return 2;
    case ING_F64: return 8;
    }
    return 0;
}

cudaError_t ingest_launch(const IngestArgs &a, int sm_count, cudaStream_t s) {
    if (a.n <= 0) return cudaSuccess;
    const int eb = ingest_element_bytes(a.fmt);
    if (eb == 0) return cudaErrorInvalidValue;
    const int vec_ok = ((reinterpret_cast<uintptr_t>(a.raw) & 15u) == 0) && ((reinterpret_cast<uintptr_t>(a.out) & 15u) == 0);
    int64_t need = ((a.n + 3) / 4 + 255) / 256;
    const int64_t cap = (int64_t)sm_count * 8;
    const int grid = (int)(need < 1 ? 1 : (need < cap ? need : cap));
    switch (a.fmt) {
    case ING_F32: ingest_kernel<ING_F32><<<grid, 256, 0, s>>>(a, vec_ok); break;
    case ING_F64: ingest_kernel<ING_F64><<<grid, 256, 0, s>>>(a, vec_ok); break;
    case ING_I16: ingest_kernel<ING_I16><<<grid, 256, 0, s>>>(a, vec_ok); break;
    case ING_I32: ingest_kernel<ING_I32><<<grid, 256, 0, s>>>(a, vec_ok); break;
    case ING_U32: ingest_kernel<ING_U32><<<grid, 256, 0, s>>>(a, vec_ok); break;
    default: return cudaErrorInvalidValue;
    }
    count_launch();
    return cudaGetLastError();
}

} // namespace afk
