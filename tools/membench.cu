// membench.cu -- read-bandwidth ceiling probes for the axis-0 reduction (exploration tool, not product code).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/membench tools/membench.cu
// Prints GB/s for (a) a linear float4 read of the whole buffer and (b) frame-strided per-pixel accumulation
// (the access pattern of reduce_kernel) for several unroll depths / block sizes / CTAs per SM.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstdint>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1);} } while (0)

__device__ __forceinline__ float4 ld_stream(const float4 *p) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ float4 ld_plain(const float4 *p) { return *p; }
__device__ __forceinline__ float4 ld_evict_first(const float4 *p) {
    float4 r;
    asm volatile("ld.global.cs.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}
struct f8 { float v[8]; };
// sm_100 256-bit load
__device__ __forceinline__ f8 ld_v8(const void *p) {
    unsigned r[8];
    asm volatile("ld.global.nc.L1::no_allocate.L2::evict_first.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]) : "l"(p));
    f8 o;
#pragma unroll
    for (int i = 0; i < 8; ++i) o.v[i] = __uint_as_float(r[i]);
    return o;
}
template <int U>
__global__ void __launch_bounds__(256) linear_read_v8(const char *p, size_t n32, float *out) {
    float acc = 0.f;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + (U - 1) * stride < n32; i += U * stride) {
        f8 v[U];
#pragma unroll
        for (int j = 0; j < U; ++j) v[j] = ld_v8(p + (i + j * stride) * 32);
#pragma unroll
        for (int j = 0; j < U; ++j)
#pragma unroll
            for (int e = 0; e < 8; ++e) acc += v[j].v[e];
    }
    if (acc == 123.456f) out[0] = acc;
}
// frame-strided, thread owns 8 pixels (32 B)
template <int BLOCK, int U, int MINB>
__global__ void __launch_bounds__(BLOCK, MINB) frames_read_v8(const char *p, int64_t n2v, int64_t nz, int64_t q, float *out) {
    const int64_t ntiles = n2v / BLOCK;
    int64_t u = (int64_t)blockIdx.x * q;
    const int64_t uend = min(u + q, ntiles * nz);
    while (u < uend) {
        const int64_t tile = u / nz, k0 = u - tile * nz;
        const int64_t k1 = min(nz, k0 + (uend - u));
        const char *ptr = p + ((k0 * n2v + tile * BLOCK + threadIdx.x) << 5);
        const int64_t fs = n2v << 5;
        float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        int64_t n = k1 - k0;
        for (; n >= U; n -= U) {
            f8 v[U];
#pragma unroll
            for (int j = 0; j < U; ++j) v[j] = ld_v8(ptr + j * fs);
            ptr += U * fs;
#pragma unroll
            for (int j = 0; j < U; ++j)
#pragma unroll
                for (int e = 0; e < 8; ++e) acc[e] += v[j].v[e];
        }
        for (; n > 0; --n) { f8 v = ld_v8(ptr); ptr += fs;
#pragma unroll
            for (int e = 0; e < 8; ++e) acc[e] += v.v[e]; }
        float *o = out + ((tile * BLOCK + threadIdx.x) << 3);
        if ((k0 == 0 && k1 == nz) || acc[0] == 123.456f) {
#pragma unroll
            for (int e = 0; e < 8; ++e) o[e] = acc[e];
        }
        u += (k1 - k0);
    }
}

__global__ void fill(float *p, size_t n) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = (float)(i & 1023) * 1e-3f;
}

template <int U, int MODE>
__global__ void __launch_bounds__(256) linear_read(const float4 *p, size_t n4, float *out) {
    float acc = 0.f;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + (U - 1) * stride < n4; i += U * stride) {
        float4 v[U];
#pragma unroll
        for (int j = 0; j < U; ++j) v[j] = MODE == 0 ? ld_stream(p + i + j * stride) : (MODE == 1 ? ld_plain(p + i + j * stride) : ld_evict_first(p + i + j * stride));
#pragma unroll
        for (int j = 0; j < U; ++j) acc += v[j].x + v[j].y + v[j].z + v[j].w;
    }
    for (; i < n4; i += stride) { float4 v = ld_stream(p + i); acc += v.x + v.y + v.z + v.w; }
    if (acc == 123.456f) out[0] = acc;
}

// contiguous chunk per CTA: CTA b reads [b*chunk, (b+1)*chunk) linearly (better DRAM page locality per CTA)
template <int U>
__global__ void __launch_bounds__(256) chunk_read(const float4 *p, size_t n4, float *out) {
    const size_t per = (n4 + gridDim.x - 1) / gridDim.x;
    const size_t b0 = (size_t)blockIdx.x * per, b1 = min(b0 + per, n4);
    float acc = 0.f;
    size_t i = b0 + threadIdx.x;
    for (; i + (U - 1) * 256 < b1; i += U * 256) {
        float4 v[U];
#pragma unroll
        for (int j = 0; j < U; ++j) v[j] = ld_stream(p + i + j * 256);
#pragma unroll
        for (int j = 0; j < U; ++j) acc += v[j].x + v[j].y + v[j].z + v[j].w;
    }
    for (; i < b1; i += 256) { float4 v = ld_stream(p + i); acc += v.x + v.y + v.z + v.w; }
    if (acc == 123.456f) out[0] = acc;
}

// frame-strided accumulation: thread owns one float4 of pixels, CTA owns a tile of BLOCK*4 pixels and a k range
template <int BLOCK, int U, int MINB>
__global__ void __launch_bounds__(BLOCK, MINB) frames_read(const float4 *p, int64_t n2v, int64_t nz, int64_t q, float4 *out) {
    // units = ntiles * nz, CTA g handles units [g*q, (g+1)*q)
    const int64_t ntiles = n2v / BLOCK;
    int64_t u = (int64_t)blockIdx.x * q;
    const int64_t uend = min(u + q, ntiles * nz);
    while (u < uend) {
        const int64_t tile = u / nz, k0 = u - tile * nz;
        const int64_t k1 = min(nz, k0 + (uend - u));
        const float4 *ptr = p + k0 * n2v + tile * BLOCK + threadIdx.x;
        float4 acc = make_float4(0, 0, 0, 0);
        int64_t n = k1 - k0;
        for (; n >= U; n -= U) {
            float4 v[U];
#pragma unroll
            for (int j = 0; j < U; ++j) v[j] = ld_stream(ptr + j * n2v);
            ptr += U * n2v;
#pragma unroll
            for (int j = 0; j < U; ++j) { acc.x += v[j].x; acc.y += v[j].y; acc.z += v[j].z; acc.w += v[j].w; }
        }
        for (; n > 0; --n) { float4 v = ld_stream(ptr); ptr += n2v; acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w; }
        if (k0 == 0 && k1 == nz) out[tile * BLOCK + threadIdx.x] = acc;
        else if (acc.x == 123.456f) out[tile * BLOCK + threadIdx.x] = acc;
        u += (k1 - k0);
    }
}

// ---- accumulation-strategy variants on the frame-strided pattern (thread owns a float4 of pixels) ----
// MODE 0: Kahan + plain p (the product kernel's arithmetic); 1: f64 accumulate + plain p; 2: pairwise-of-U then Kahan + p;
// 3: f64 only; 4: f64 of pairwise-of-U batch sums + p
template <int U, int MODE, int MINB>
__global__ void __launch_bounds__(256, MINB) frames_acc(const float4 *p, int64_t n2v, int64_t nz, int64_t q, float4 *out) {
    const int64_t ntiles = n2v / 256;
    int64_t u = (int64_t)blockIdx.x * q;
    const int64_t uend = min(u + q, ntiles * nz);
    while (u < uend) {
        const int64_t tile = u / nz, k0 = u - tile * nz;
        const int64_t k1 = min(nz, k0 + (uend - u));
        const float4 *ptr = p + k0 * n2v + tile * 256 + threadIdx.x;
        float s[4] = {-0.f, -0.f, -0.f, -0.f}, c[4] = {0, 0, 0, 0}, pl[4] = {-0.f, -0.f, -0.f, -0.f};
        double d[4] = {0, 0, 0, 0};
        int64_t n = k1 - k0;
        for (; n >= U; n -= U) {
            float4 v[U];
#pragma unroll
            for (int j = 0; j < U; ++j) v[j] = ld_stream(ptr + j * n2v);
            ptr += U * n2v;
            float e[U][4];
#pragma unroll
            for (int j = 0; j < U; ++j) { e[j][0] = v[j].x; e[j][1] = v[j].y; e[j][2] = v[j].z; e[j][3] = v[j].w; }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                if (MODE != 3) {
#pragma unroll
                    for (int j = 0; j < U; ++j) pl[i] = __fadd_rn(pl[i], e[j][i]);
                }
                if (MODE == 0) {
#pragma unroll
                    for (int j = 0; j < U; ++j) {
                        float y = __fsub_rn(e[j][i], c[i]);
                        float t = __fadd_rn(s[i], y);
                        c[i] = __fsub_rn(__fsub_rn(t, s[i]), y);
                        s[i] = t;
                    }
                } else if (MODE == 1 || MODE == 3) {
#pragma unroll
                    for (int j = 0; j < U; ++j) d[i] += (double)e[j][i];
                } else {
                    float t[U];
#pragma unroll
                    for (int j = 0; j < U; ++j) t[j] = e[j][i];
#pragma unroll
                    for (int w = U / 2; w >= 1; w >>= 1)
#pragma unroll
                        for (int j = 0; j < w; ++j) t[j] = __fadd_rn(t[j], t[j + w]);
                    if (MODE == 2) {
                        float y = __fsub_rn(t[0], c[i]);
                        float tt = __fadd_rn(s[i], y);
                        c[i] = __fsub_rn(__fsub_rn(tt, s[i]), y);
                        s[i] = tt;
                    } else {
                        d[i] += (double)t[0];
                    }
                }
            }
        }
        float4 r;
        if (MODE == 0 || MODE == 2) r = make_float4(s[0] - c[0] + pl[0] * 1e-30f, s[1] - c[1] + pl[1] * 1e-30f, s[2] - c[2] + pl[2] * 1e-30f, s[3] - c[3] + pl[3] * 1e-30f);
        else r = make_float4((float)d[0] + pl[0] * 1e-30f, (float)d[1] + pl[1] * 1e-30f, (float)d[2] + pl[2] * 1e-30f, (float)d[3] + pl[3] * 1e-30f);
        if ((k0 == 0 && k1 == nz) || r.x == 123.456f) out[tile * 256 + threadIdx.x] = r;
        u += (k1 - k0);
    }
}

template <typename F> float time_it(F f, int reps = 20) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int i = 0; i < 3; ++i) f();
    CK(cudaDeviceSynchronize());
    float best = 1e9f, tot = 0;
    for (int i = 0; i < reps; ++i) {
        cudaEventRecord(e0); f(); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); best = ms < best ? ms : best; tot += ms;
    }
    CK(cudaGetLastError());
    printf("   best %.4f ms avg %.4f ms", best, tot / reps);
    return tot / reps;
}

int main() {
    const int64_t nz = 3840, n2 = 1024 * 1024;
    const size_t n = (size_t)nz * n2;
    float *buf, *out;
    CK(cudaMalloc(&buf, n * 4));
    CK(cudaMalloc(&out, n2 * 4 * 2));
    fill<<<148 * 16, 256>>>(buf, n);
    CK(cudaDeviceSynchronize());
    const double gb = n * 4 / 1e9;
    const size_t n4 = n / 4;
#define RUN(label, ...) { printf("%-46s", label); float ms = time_it([&] { __VA_ARGS__; }); printf("  %8.1f GB/s\n", gb / (ms * 1e-3)); }
    for (int ctas : {148 * 2, 148 * 4, 148 * 8, 148 * 16}) {
        char l[96];
        snprintf(l, 96, "linear U4 nc.noalloc grid %d", ctas); RUN(l, (linear_read<4, 0><<<ctas, 256>>>((float4 *)buf, n4, out)));
        snprintf(l, 96, "linear U8 nc.noalloc grid %d", ctas); RUN(l, (linear_read<8, 0><<<ctas, 256>>>((float4 *)buf, n4, out)));
        snprintf(l, 96, "linear U16 nc.noalloc grid %d", ctas); RUN(l, (linear_read<16, 0><<<ctas, 256>>>((float4 *)buf, n4, out)));
    }
    RUN("linear U8 plain ld grid 1184", (linear_read<8, 1><<<1184, 256>>>((float4 *)buf, n4, out)));
    RUN("linear U8 ld.cs grid 1184", (linear_read<8, 2><<<1184, 256>>>((float4 *)buf, n4, out)));
    for (int ctas : {148 * 4, 148 * 8, 148 * 32, 148 * 128}) {
        char l[96];
        snprintf(l, 96, "chunk U8 grid %d", ctas); RUN(l, (chunk_read<8><<<ctas, 256>>>((float4 *)buf, n4, out)));
    }
    RUN("linear v8 U4 grid 592", (linear_read_v8<4><<<592, 256>>>((const char *)buf, n4 / 2, out)));
    RUN("linear v8 U4 grid 1184", (linear_read_v8<4><<<1184, 256>>>((const char *)buf, n4 / 2, out)));
    RUN("linear v8 U8 grid 592", (linear_read_v8<8><<<592, 256>>>((const char *)buf, n4 / 2, out)));
    {
        const int64_t n2v8 = n2 / 8;
        auto fr8 = [&](auto kern, int block, int ctas_per_sm, const char *name) {
            const int64_t ntiles = n2v8 / block, units = ntiles * nz;
            const int64_t G = 148 * ctas_per_sm;
            const int64_t q = (units + G - 1) / G;
            const int grid = (int)((units + q - 1) / q);
            char l[96];
            snprintf(l, 96, "%s grid %d", name, grid);
            RUN(l, (kern<<<grid, block>>>((const char *)buf, n2v8, nz, q, out)));
        };
        fr8(frames_read_v8<256, 4, 4>, 256, 4, "frames v8 B256 U4 4/SM");
        fr8(frames_read_v8<256, 4, 3>, 256, 3, "frames v8 B256 U4 3/SM");
        fr8(frames_read_v8<256, 8, 2>, 256, 2, "frames v8 B256 U8 2/SM");
        fr8(frames_read_v8<128, 4, 8>, 128, 8, "frames v8 B128 U4 8/SM");
        fr8(frames_read_v8<128, 8, 4>, 128, 4, "frames v8 B128 U8 4/SM");
        fr8(frames_read_v8<512, 4, 2>, 512, 2, "frames v8 B512 U4 2/SM");
    }
    const int64_t n2v = n2 / 4;
    auto fr = [&](auto kern, int block, int ctas_per_sm, const char *name) {
        const int64_t ntiles = n2v / block, units = ntiles * nz;
        const int64_t G = 148 * ctas_per_sm;
        const int64_t q = (units + G - 1) / G;
        const int grid = (int)((units + q - 1) / q);
        char l[96];
        snprintf(l, 96, "%s grid %d", name, grid);
        RUN(l, (kern<<<grid, block>>>((float4 *)buf, n2v, nz, q, (float4 *)out)));
    };
    fr(frames_acc<8, 0, 4>, 256, 4, "acc kahan+p U8 4/SM");
    fr(frames_acc<4, 0, 4>, 256, 4, "acc kahan+p U4 4/SM");
    fr(frames_acc<4, 0, 6>, 256, 6, "acc kahan+p U4 6/SM");
    fr(frames_acc<8, 1, 4>, 256, 4, "acc f64+p U8 4/SM");
    fr(frames_acc<4, 1, 4>, 256, 4, "acc f64+p U4 4/SM");
    fr(frames_acc<4, 1, 6>, 256, 6, "acc f64+p U4 6/SM");
    fr(frames_acc<8, 3, 4>, 256, 4, "acc f64 only U8 4/SM");
    fr(frames_acc<8, 2, 4>, 256, 4, "acc pair8->kahan+p U8 4/SM");
    fr(frames_acc<4, 2, 4>, 256, 4, "acc pair4->kahan+p U4 4/SM");
    fr(frames_acc<4, 2, 6>, 256, 6, "acc pair4->kahan+p U4 6/SM");
    fr(frames_acc<8, 4, 4>, 256, 4, "acc pair8->f64+p U8 4/SM");
    fr(frames_acc<4, 4, 4>, 256, 4, "acc pair4->f64+p U4 4/SM");
    fr(frames_acc<4, 4, 6>, 256, 6, "acc pair4->f64+p U4 6/SM");
    fr(frames_read<256, 4, 4>, 256, 4, "frames B256 U4 4/SM");
    fr(frames_read<256, 8, 4>, 256, 4, "frames B256 U8 4/SM");
    fr(frames_read<256, 16, 2>, 256, 2, "frames B256 U16 2/SM");
    fr(frames_read<256, 8, 6>, 256, 6, "frames B256 U8 6/SM");
    fr(frames_read<256, 4, 8>, 256, 8, "frames B256 U4 8/SM");
    fr(frames_read<256, 8, 8>, 256, 8, "frames B256 U8 8/SM");
    fr(frames_read<128, 8, 8>, 128, 8, "frames B128 U8 8/SM");
    fr(frames_read<128, 8, 16>, 128, 16, "frames B128 U8 16/SM");
    fr(frames_read<512, 8, 2>, 512, 2, "frames B512 U8 2/SM");
    fr(frames_read<512, 8, 4>, 512, 4, "frames B512 U8 4/SM");
    fr(frames_read<1024, 8, 1>, 1024, 1, "frames B1024 U8 1/SM");
    fr(frames_read<1024, 8, 2>, 1024, 2, "frames B1024 U8 2/SM");
    fr(frames_read<1024, 4, 2>, 1024, 2, "frames B1024 U4 2/SM");
    RUN("cudaMemcpy D2D 8GB (read+write bytes)", CK(cudaMemcpyAsync(buf, buf + n / 2, n * 2, cudaMemcpyDeviceToDevice)));
    return 0;
}
