// c1_probe.cu -- where do the 33 us of the MaNGA-shape integral (74x74x4563 f32, 100 MB) go?  Exploration tool, not product code.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/c1_probe tools/c1_probe.cu
// Probes, every one timed with CUDA events (single launch between two events, median of 50; a 256 MiB buffer is
// overwritten before each launch so that L2 is cold):
//   null        an empty kernel: the event-to-event launch floor
//   linear      plain float4 grid-stride read of the 100 MB (no structure at all)
//   bulk_chunk  cp.async.bulk ring, every CTA streams one contiguous chunk (no pixel structure): the TMA read floor
//   cluster     the candidate kernel: one cluster per pixel tile, the frames split over the cluster's CTAs, bulk-copy ring,
//               merge through distributed shared memory; checked against an f64 sum on the host
#include <cooperative_groups.h>
#include <cuda_runtime.h>
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
namespace cg = cooperative_groups;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1);} } while (0)

// ---- PTX wrappers ----
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *b, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *b, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *b) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *b, uint32_t parity) {
    asm volatile(
        "{\n .reg .pred p;\n WAIT_%=:\n mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n @p bra DONE_%=;\n bra WAIT_%=;\n DONE_%=:\n}\n" ::"r"(smem_u32(b)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src),
                 "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ float4 ld_stream(const float4 *p) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}

__global__ void null_kernel(float *out) {
    if (out == nullptr && threadIdx.x == 12345) printf("x");
}

__global__ void fill(float *p, size_t n, unsigned seed) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        unsigned h = (unsigned)i * 2654435761u + seed;
        h ^= h >> 15; h *= 2246822519u; h ^= h >> 13;
        p[i] = (float)(h & 0xFFFF) * (1.0f / 65536.0f) + 0.25f * (float)((h >> 16) & 7);
    }
}
__global__ void flush_kernel(float4 *p, size_t n4) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) p[i] = make_float4(1, 2, 3, 4);
}

template <int U>
__global__ void __launch_bounds__(256) linear_read(const float4 *p, size_t n4, float *out) {
    float acc = 0.f;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + (U - 1) * stride < n4; i += U * stride) {
        float4 v[U];
#pragma unroll
        for (int j = 0; j < U; ++j) v[j] = ld_stream(p + i + j * stride);
#pragma unroll
        for (int j = 0; j < U; ++j) acc += v[j].x + v[j].y + v[j].z + v[j].w;
    }
    for (; i < n4; i += stride) { float4 v = ld_stream(p + i); acc += v.x + v.y + v.z + v.w; }
    if (acc == 123.456f) out[0] = acc;
}

// every CTA streams one contiguous chunk through a ring of `stages` buffers of `piece` bytes (one bulk copy each)
__global__ void __launch_bounds__(288, 1) bulk_chunk_kernel(const char *p, size_t bytes, int piece, int stages, float *out) {
    extern __shared__ __align__(128) unsigned char smem[];
    uint64_t *full = reinterpret_cast<uint64_t *>(smem);
    uint64_t *empty = full + stages;
    unsigned char *ring = smem + 1024;
    const size_t per = ((bytes + gridDim.x - 1) / gridDim.x + 15) / 16 * 16;
    const size_t b0 = (size_t)blockIdx.x * per, b1 = min(b0 + per, bytes);
    const int npieces = b0 < b1 ? (int)((b1 - b0 + piece - 1) / piece) : 0;
    const int tid = threadIdx.x, warp = tid >> 5;
    if (tid == 0) {
        for (int s = 0; s < stages; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, 8); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (warp == 8) {
        if (tid == 256) {
            for (int i = 0; i < npieces; ++i) {
                const int s = i % stages, it = i / stages;
                if (it > 0) mbar_wait(empty + s, (it - 1) & 1);
                const size_t off = b0 + (size_t)i * piece;
                const uint32_t nb = (uint32_t)min((size_t)piece, b1 - off);
                mbar_expect_tx(full + s, nb);
                bulk_g2s(ring + (size_t)s * piece, p + off, nb, full + s);
            }
        }
    } else {
        float acc = 0.f;
        for (int i = 0; i < npieces; ++i) {
            const int s = i % stages, it = i / stages;
            mbar_wait(full + s, it & 1);
            const size_t off = b0 + (size_t)i * piece;
            const int n4 = (int)(min((size_t)piece, b1 - off) / 16);
            const float4 *src = reinterpret_cast<const float4 *>(ring + (size_t)s * piece);
            for (int j = tid; j < n4; j += 256) { float4 v = src[j]; acc += v.x + v.y + v.z + v.w; }
            __syncwarp();
            if ((tid & 31) == 0) mbar_arrive(empty + s);
        }
        if (acc == 123.456f) out[0] = acc;
    }
}

// ---- candidate: one cluster per pixel tile, frames split over the cluster, bulk-copy ring, DSMEM merge ----
struct CArgs {
    const float *cube;
    int64_t n2;
    int nwin;
    int64_t start[3];
    int64_t woff[4];
    int64_t ktot;
    int tile4; // float4 lanes per tile
    int F, S, J;
    float *out;
};
struct Acc { float s, c, p; };
__device__ __forceinline__ void acc_add(Acc &a, float x) {
    a.p = __fadd_rn(a.p, x);
    float y = __fsub_rn(x, a.c);
    float t = __fadd_rn(a.s, y);
    a.c = __fsub_rn(__fsub_rn(t, a.s), y);
    a.s = t;
}

__global__ void __launch_bounds__(288, 1) cluster_reduce(const CArgs a) {
    extern __shared__ __align__(128) unsigned char smem[];
    cg::cluster_group cl = cg::this_cluster();
    const int C = (int)cl.num_blocks(), rank = (int)cl.block_rank();
    const int tile = blockIdx.x / C;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int64_t l0 = (int64_t)tile * a.tile4;
    const int n4 = (int)min((int64_t)a.tile4, a.n2 / 4 - l0);
    const uint32_t pb = (uint32_t)n4 * 16u;
    const int tpx = a.tile4 * 4;

    uint64_t *full = reinterpret_cast<uint64_t *>(smem);
    uint64_t *empty = full + a.S;
    double2 *part = reinterpret_cast<double2 *>(smem + 1024);              // [3][tpx]
    double2 *comb = part + 3 * tpx;                                         // [J][tpx]
    float4 *ring = reinterpret_cast<float4 *>(comb + (size_t)a.J * tpx);    // [S][F][tile4]

    const int64_t kc = (a.ktot + C - 1) / C;
    const int64_t r0 = (int64_t)rank * kc, r1 = min(a.ktot, r0 + kc);

    if (tid == 0) {
        for (int s = 0; s < a.S; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, 8); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int i = tid; i < 3 * tpx; i += blockDim.x) part[i] = make_double2(-0.0, -0.0);
    __syncthreads();

    if (warp == 8) {
        int i = 0;
        for (int w = 0; w < a.nwin; ++w) {
            const int64_t lo = max(r0, a.woff[w]), hi = min(r1, a.woff[w + 1]);
            for (int64_t k = lo; k < hi; k += a.F, ++i) {
                const int nf = (int)min((int64_t)a.F, hi - k);
                const int s = i % a.S, it = i / a.S;
                if (lane == 0) {
                    if (it > 0) mbar_wait(empty + s, (it - 1) & 1);
                    mbar_expect_tx(full + s, (uint32_t)nf * pb);
                }
                __syncwarp();
                if (lane < nf)
                    bulk_g2s(ring + ((size_t)s * a.F + lane) * a.tile4, a.cube + (a.start[w] + (k - a.woff[w]) + lane) * a.n2 + l0 * 4, pb, full + s);
            }
        }
    } else {
        const int li = tid % a.tile4, j = tid / a.tile4;
        const bool active = (j < a.J) && (li < n4);
        int i = 0;
        for (int w = 0; w < a.nwin; ++w) {
            const int64_t lo = max(r0, a.woff[w]), hi = min(r1, a.woff[w + 1]);
            if (lo >= hi) continue;
            Acc acc[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) { acc[e].s = -0.0f; acc[e].c = 0.0f; acc[e].p = -0.0f; }
            for (int64_t k = lo; k < hi; k += a.F, ++i) {
                const int nf = (int)min((int64_t)a.F, hi - k);
                const int s = i % a.S, it = i / a.S;
                mbar_wait(full + s, it & 1);
                if (active) {
                    const float4 *src = ring + (size_t)s * a.F * a.tile4 + li;
                    for (int f = j; f < nf; f += a.J) {
                        const float4 v = src[(size_t)f * a.tile4];
                        acc_add(acc[0], v.x); acc_add(acc[1], v.y); acc_add(acc[2], v.z); acc_add(acc[3], v.w);
                    }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(empty + s);
            }
            if (active) {
#pragma unroll
                for (int e = 0; e < 4; ++e) comb[(size_t)j * tpx + li * 4 + e] = make_double2((double)acc[e].s - (double)acc[e].c, (double)acc[e].p);
            }
            asm volatile("bar.sync 1, 256;" ::: "memory");
            if (active && j == 0) {
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    double d = -0.0;
                    float pl = -0.0f;
                    for (int jj = 0; jj < a.J; ++jj) {
                        const double2 v = comb[(size_t)jj * tpx + li * 4 + e];
                        d += v.x;
                        pl = __fadd_rn(pl, (float)v.y);
                    }
                    part[(size_t)w * tpx + li * 4 + e] = make_double2(d, (double)pl);
                }
            }
            asm volatile("bar.sync 1, 256;" ::: "memory");
        }
    }
    cl.sync();
    {
        const int npx = n4 * 4;
        const int pc = (npx + C - 1) / C;
        const int q = rank * pc + tid;
        if (tid < pc && q < npx) {
            float S[3];
            for (int w = 0; w < a.nwin; ++w) {
                double d = -0.0;
                float pl = -0.0f;
                for (int cc = 0; cc < C; ++cc) {
                    const double2 *rp = cl.map_shared_rank(part, cc);
                    const double2 v = rp[(size_t)w * tpx + q];
                    d += v.x;
                    pl = __fadd_rn(pl, (float)v.y);
                }
                float r = (float)d;
                if (!(fabsf(pl) <= 3.402823466e38f)) r = pl;
                S[w] = r;
            }
            a.out[l0 * 4 + q] = S[0];
        }
    }
    cl.sync();
}

// ---- host ----
static float *g_flush = nullptr;
static const size_t kFlushBytes = 256ull << 20;
template <typename L> static void time_it(const char *what, double bytes, L launch) {
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    for (int i = 0; i < 5; ++i) launch();
    CK(cudaDeviceSynchronize());
    std::vector<float> cold, warm;
    for (int i = 0; i < 40; ++i) {
        flush_kernel<<<1184, 256>>>(reinterpret_cast<float4 *>(g_flush), kFlushBytes / 16);
        CK(cudaEventRecord(e0));
        launch();
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        cold.push_back(ms * 1e3f);
    }
    for (int i = 0; i < 40; ++i) {
        CK(cudaEventRecord(e0));
        launch();
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        warm.push_back(ms * 1e3f);
    }
    float b2b = 1e30f;
    for (int r = 0; r < 5; ++r) {
        CK(cudaEventRecord(e0));
        for (int i = 0; i < 20; ++i) launch();
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        b2b = std::min(b2b, ms * 1e3f / 20);
    }
    std::sort(cold.begin(), cold.end());
    std::sort(warm.begin(), warm.end());
    const float c = cold[cold.size() / 2], w = warm[warm.size() / 2];
    printf("%-58s cold %6.2f us (min %6.2f)  warm %6.2f us  b2b %6.2f us   %7.1f GB/s cold\n", what, c, cold[0], w, b2b, bytes / c * 1e-3);
    fflush(stdout);
    CK(cudaEventDestroy(e0));
    CK(cudaEventDestroy(e1));
}

int main(int argc, char **argv) {
    int nx = 74, ny = 74, nz = 4563;
    if (argc > 3) { nz = atoi(argv[1]); ny = atoi(argv[2]); nx = atoi(argv[3]); }
    const int64_t n2 = (int64_t)ny * nx;
    const size_t n = (size_t)nz * n2;
    const double bytes = (double)n * 4;
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    printf("cube %dx%dx%d  %.1f MB  SMs %d\n", nz, ny, nx, bytes / 1e6, sms);
    float *cube, *out;
    CK(cudaMalloc(&cube, n * 4));
    CK(cudaMalloc(&out, (size_t)n2 * 4 + 4096));
    CK(cudaMalloc(&g_flush, kFlushBytes));
    fill<<<1184, 256>>>(cube, n, 12345u);
    CK(cudaDeviceSynchronize());
    std::vector<float> h(n);
    CK(cudaMemcpy(h.data(), cube, n * 4, cudaMemcpyDeviceToHost));
    std::vector<double> want(n2, 0.0);
    for (int64_t k = 0; k < nz; ++k)
        for (int64_t p = 0; p < n2; ++p) want[p] += (double)h[(size_t)k * n2 + p];

    time_it("null kernel", bytes, [&] { null_kernel<<<1, 32>>>(out); });
    for (int g : {148, 296, 592, 1184})
        { char nm[64]; snprintf(nm, 64, "linear float4 U8 grid %d", g); time_it(nm, bytes, [&] { linear_read<8><<<g, 256>>>(reinterpret_cast<const float4 *>(cube), n / 4, out); }); }
    CK(cudaFuncSetAttribute(bulk_chunk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
    for (int piece : {4096, 8192, 16384, 32768})
        for (int stages : {4, 8, 12}) {
            if ((size_t)piece * stages > 200 * 1024) continue;
            char nm[96];
            snprintf(nm, 96, "bulk_chunk grid %d piece %d stages %d (%d KB in flight)", sms, piece, stages, piece * stages / 1024);
            time_it(nm, bytes, [&] { bulk_chunk_kernel<<<sms, 288, 1024 + piece * stages>>>(reinterpret_cast<const char *>(cube), n * 4, piece, stages, out); });
        }

    CK(cudaFuncSetAttribute(cluster_reduce, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
    CK(cudaFuncSetAttribute(cluster_reduce, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    for (int C : {2, 4, 8, 16}) {
        // how many clusters of this size are co-resident?
        cudaLaunchConfig_t qc = {};
        qc.gridDim = dim3(C * 64);
        qc.blockDim = dim3(288);
        qc.dynamicSmemBytes = 200 * 1024;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = C; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        qc.attrs = at; qc.numAttrs = 1;
        int maxcl = 0;
        cudaError_t qe = cudaOccupancyMaxActiveClusters(&maxcl, cluster_reduce, &qc);
        printf("cluster size %d: max active clusters %d (%s) -> %d CTAs\n", C, maxcl, cudaGetErrorString(qe), maxcl * C);
        if (qe != cudaSuccess || maxcl < 1) { cudaGetLastError(); continue; }
        const int64_t n4tot = n2 / 4;
        for (int ncl : {maxcl, maxcl - 1, maxcl * 3 / 4}) {
            if (ncl < 1) continue;
            int tile4 = (int)((n4tot + ncl - 1) / ncl);
            if (tile4 > 256) continue;
            const int ntiles = (int)((n4tot + tile4 - 1) / tile4);
            const int J = std::min(8, 256 / tile4);
            for (int F : {8, 16, 32}) {
                CArgs a{};
                a.cube = cube; a.n2 = n2; a.nwin = 1; a.start[0] = 0; a.woff[0] = 0; a.woff[1] = a.woff[2] = a.woff[3] = nz; a.ktot = nz;
                a.tile4 = tile4; a.F = F; a.J = J; a.out = out;
                const size_t fixed = 1024 + (size_t)(3 + J) * tile4 * 4 * 16;
                const size_t stage = (size_t)F * tile4 * 16;
                int S = (int)((200 * 1024 - fixed) / stage);
                if (S > 16) S = 16;
                if (S < 2) continue;
                a.S = S;
                cudaLaunchConfig_t lc = {};
                lc.gridDim = dim3(ntiles * C);
                lc.blockDim = dim3(288);
                lc.dynamicSmemBytes = fixed + stage * S;
                lc.attrs = at; lc.numAttrs = 1;
                CK(cudaMemset(out, 0, (size_t)n2 * 4));
                cudaError_t le = cudaLaunchKernelEx(&lc, cluster_reduce, a);
                if (le != cudaSuccess) { printf("  launch failed: %s\n", cudaGetErrorString(le)); cudaGetLastError(); continue; }
                CK(cudaDeviceSynchronize());
                std::vector<float> got(n2);
                CK(cudaMemcpy(got.data(), out, (size_t)n2 * 4, cudaMemcpyDeviceToHost));
                double worst = 0;
                for (int64_t p = 0; p < n2; ++p) worst = std::max(worst, std::fabs((double)got[p] - want[p]) / std::fabs(want[p]));
                char nm[128];
                snprintf(nm, 128, "cluster C%d tiles %d tile4 %d J %d F %d S %d (%zu KB ring) err %.1e", C, ntiles, tile4, J, F, S, stage * S / 1024, worst);
                time_it(nm, bytes, [&] { cudaLaunchKernelEx(&lc, cluster_reduce, a); });
            }
        }
    }
    CK(cudaDeviceSynchronize());
    return 0;
}
