// noalloc.cu -- afk_k_reduce_static: the axis-0 windowed reduction (afk_k_reduce) for callers that cannot allocate.
//
// No counterpart in the reference.  Some callers cannot call cudaMalloc: code
// under CUDA graph capture, tight serving loops, harnesses that forbid it.  afk_k_reduce takes its split-K scratch (f64
// partials of pixel tiles shared between CTAs + per-tile arrival counters) from the stream-ordered pool; this entry point takes
// it from a file-scope __device__ array instead and calls no cudaMalloc / cudaFree / allocating library routine, ever.
//
// * The array is part of the module image: every GPU that loads the library holds kStaticScratchBytes (32.25 MiB) until exit.
// * A problem whose plan needs more than kStaticPartialBytes of partials or kStaticCounters counters is rejected with
//   AFK_ERR_UNEXPECTED_INPUT and a message that states both figures (use afk_k_reduce for those).
// * Every call on a device shares the one array: calls that overlap on the GPU (different streams, or one CUDA graph launched
//   twice without a dependency) race on the partials and the counters.  Serialise them.
// * The array keeps the last call's contents.  The partials are written before they are read; the arrival counters must start
//   at zero and are cleared by a cudaMemsetAsync on the caller's stream in every call (a memset node under capture).
// * In host code the array's NAME is a host-side stand-in, not a device address.  The address is taken once per device, with
//   that device current, by cudaGetSymbolAddress (below), and that pointer is what the kernel is given.
//
// Same arithmetic, same kernels, same launch plan as afk_k_reduce: the results are bit-identical.
#include <mutex>

#include "host.h"

namespace afk {

constexpr size_t kStaticPartialBytes = (size_t)32 << 20; // split-K partials + per-CTA range slots
constexpr size_t kStaticCounters = (size_t)64 << 10;     // per-tile arrival counters (+ 1): images up to 64 Mi pixels
constexpr size_t kStaticScratchBytes = kStaticPartialBytes + kStaticCounters * sizeof(unsigned);

static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

static std::mutex g_static_mu;
static unsigned char *g_static_ptr[16] = {}; // device address of g_static_scratch on each device, once known

// set-up, once per device, with that device current
static Error static_scratch(int device, unsigned char **out) {
    if (device < 0 || device >= 16) return make_error(AFK_ERR_CUDA, "CUDA: device index out of range");
    std::lock_guard<std::mutex> lk(g_static_mu);
    if (!g_static_ptr[device]) {
        int cur = -1;
        AFK_CUDA(cudaGetDevice(&cur));
        if (cur != device) return make_error(AFK_ERR_CUDA, "CUDA: the static scratch is resolved with its device current");
        unsigned char *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void **>(&p), g_static_scratch);
        if (e != cudaSuccess) return cuda_error(e, "cudaGetSymbolAddress(g_static_scratch)");
        if (!p || (reinterpret_cast<uintptr_t>(p) & 255u) != 0) return make_error(AFK_ERR_CUDA, "CUDA: bad address for the static scratch");
        g_static_ptr[device] = p;
    }
    *out = g_static_ptr[device];
    return ok();
}

} // namespace afk

using namespace afk;

extern "C" {

afk_status afk_k_reduce_static(const float *cube, int64_t nz, int64_t n2, int nwin, const int64_t *starts, const int64_t *ends, int epilogue,
                               const float *params, int sum_mode, float *out, void *cuda_stream) {
    // the checks of afk_k_reduce (ops.cpp), same wording
    if (!cube || !out || !starts || !ends) return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "null argument"));
    if (nwin < 1 || nwin > kMaxWindows) return fail(unexpected(fmt("between 1 and %d windows are supported, got %d", kMaxWindows, nwin)));
    const int need = epilogue == EPI_EW ? 3 : (epilogue == EPI_MEAN_RATIO ? 2 : 1);
    if (epilogue < EPI_SUM || epilogue > EPI_EW || nwin != need) return fail(unexpected("epilogue and window count do not match"));
    ReduceArgs a{};
    a.cube = cube;
    a.n2 = n2;
    a.nwin = nwin;
    for (int w = 0; w < nwin; ++w) {
        if (starts[w] < 0) return fail(try_into_unsigned("start", starts[w]));
        if (!(starts[w] < ends[w])) return fail(unexpected(fmt("Expected start < end, but got %lld < %lld", (long long)starts[w], (long long)ends[w])));
        if (!(ends[w] < nz + 1))
            return fail(unexpected(fmt("'end' greater or equal to the input image 'image''s frame count (expected %lld < %lld)",
                                       (long long)ends[w], (long long)(nz + 1))));
        a.start[w] = starts[w];
        a.end[w] = ends[w];
    }
    a.epilogue = epilogue;
    a.mode = sum_mode;
    if (epilogue == EPI_EW && !params) return fail(unexpected("the equivalent-width epilogue needs params[5] = {a, b, fl, max, sign}"));
    for (int i = 0; i < 4; ++i) a.p[i] = epilogue == EPI_EW ? params[i] : 0.0f;
    a.ew_sign = epilogue == EPI_EW ? params[4] : 1.0f;
    a.out = out;
    if (n2 <= 0) return AFK_OK;

    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess) {
        cudaGetLastError();
        return fail(make_error(AFK_ERR_CUDA, "CUDA: no CUDA device available (this library has no CPU fallback)"));
    }
    const ReducePlan plan = reduce_plan(a, sm_count(dev));
    if (plan.partial_bytes > kStaticPartialBytes || plan.n_counters > kStaticCounters)
        return fail(unexpected(fmt("afk_k_reduce_static: this problem needs %zu bytes of split-K scratch and %zu arrival counters, the static scratch "
                                   "holds %zu bytes and %zu counters (use afk_k_reduce, which allocates)",
                                   plan.partial_bytes, plan.n_counters, kStaticPartialBytes, kStaticCounters)));
    unsigned char *base = nullptr;
    if (Error e = static_scratch(dev, &base)) return fail(e);
    ReduceWorkspace ws;
    ws.partials = base;
    ws.partial_bytes = kStaticPartialBytes;
    ws.counters = reinterpret_cast<unsigned *>(base + kStaticPartialBytes);
    ws.n_counters = kStaticCounters;
    cudaStream_t s = static_cast<cudaStream_t>(cuda_stream);
    // the counters must start at zero; the array keeps whatever the last call (or a call that failed half-way) left
    cudaError_t ce = cudaMemsetAsync(ws.counters, 0, plan.n_counters * sizeof(unsigned), s);
    if (ce == cudaSuccess) ce = reduce_launch(a, plan, ws, s);
    if (ce != cudaSuccess) return fail(cuda_error(ce, "afk_k_reduce_static launch"));
    return AFK_OK;
}

} // extern "C"
