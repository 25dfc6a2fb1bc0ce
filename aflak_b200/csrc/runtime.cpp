// runtime.cpp -- execution context, device-resident image handles and the IOValue model.
// Reference citations are relative to /root/reference/src/.
#include <chrono>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <unordered_map>

#include "host.h"

namespace afk {

bool trace_enabled();

// ---- errors ------------------------------------------------------------------------------------------
std::string fmt(const char *f, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, f);
    vsnprintf(buf, sizeof buf, f, ap);
    va_end(ap);
    return std::string(buf);
}

Error make_error(afk_status st, const std::string &msg) {
    Error e;
    e.status = st;
    e.message = msg;
    return e;
}

Error cuda_error(cudaError_t e, const char *what) {
    // clear the sticky-less error state so that later calls on this thread start clean
    cudaGetLastError();
    return make_error(AFK_ERR_CUDA, fmt("CUDA: %s (%s) in %s", cudaGetErrorString(e), cudaGetErrorName(e), what));
}

ThreadCtx &tls() {
    static thread_local ThreadCtx ctx;
    return ctx;
}

static std::atomic<int> g_default_sum_mode{SUM_COMPENSATED};
int effective_sum_mode() {
    const int m = tls().sum_mode;
    return m == SUM_INHERIT ? g_default_sum_mode.load(std::memory_order_relaxed) : m;
}

void set_last_error(const std::string &msg) { tls().last_error = msg; }

afk_status fail(const Error &e) {
    if (e.status != AFK_OK) set_last_error(e.message);
    return e.status;
}

// ---- devices / streams ---------------------------------------------------------------------------------
static std::mutex g_dev_mu;
static int g_sm_count[16] = {0};
static bool g_pool_configured[16] = {false};
// Cubes get a memory pool of their own.  In the default pool the small stream-ordered allocations of the operators
// (2-D results, split-K partials, flags) carve pieces out of a freed 16 GB block, and the next cube upload then makes
// the driver stitch a new 16 GB virtual range together page by page: measured 100-500 ms per cudaMallocAsync on B200
// (profiles/r1_e2e_alloc_trace.txt) against 0.01 ms for an exact-fit reuse.
static cudaMemPool_t g_big_pool[16] = {nullptr};
constexpr size_t kBigAllocBytes = (size_t)32 << 20;

int sm_count(int device) {
    if (device < 0 || device >= 16) return 148;
    std::lock_guard<std::mutex> lk(g_dev_mu);
    if (!g_sm_count[device]) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || n <= 0) n = 148;
        g_sm_count[device] = n;
    }
    return g_sm_count[device];
}

static Error configure_pool(int device) {
    std::lock_guard<std::mutex> lk(g_dev_mu);
    if (device >= 0 && device < 16 && !g_pool_configured[device]) {
        cudaMemPool_t pool;
        AFK_CUDA(cudaDeviceGetDefaultMemPool(&pool, device));
        uint64_t threshold = UINT64_MAX; // keep freed blocks cached: chained DAG nodes reuse them without cudaMalloc
        AFK_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &threshold));
        cudaMemPoolProps props = {};
        props.allocType = cudaMemAllocationTypePinned;
        props.handleTypes = cudaMemHandleTypeNone;
        props.location.type = cudaMemLocationTypeDevice;
        props.location.id = device;
        AFK_CUDA(cudaMemPoolCreate(&g_big_pool[device], &props));
        AFK_CUDA(cudaMemPoolSetAttribute(g_big_pool[device], cudaMemPoolAttrReleaseThreshold, &threshold));
        g_pool_configured[device] = true;
    }
    return ok();
}

Error current_stream(cudaStream_t *out, int *device) {
    ThreadCtx &c = tls();
    if (c.device < 0) {
        int n = 0;
        cudaError_t e = cudaGetDeviceCount(&n);
        if (e != cudaSuccess || n <= 0) {
            cudaGetLastError();
            return make_error(AFK_ERR_CUDA, "CUDA: no CUDA device available (this library has no CPU fallback)");
        }
        int cur = 0;
        AFK_CUDA(cudaGetDevice(&cur));
        c.device = cur;
    }
    AFK_CUDA(cudaSetDevice(c.device));
    if (Error e = configure_pool(c.device)) return e;
    if (c.external) {
        *out = c.external;
    } else {
        if (c.device >= 16) return make_error(AFK_ERR_CUDA, "CUDA: device index out of range");
        if (!c.own[c.device]) AFK_CUDA(cudaStreamCreateWithFlags(&c.own[c.device], cudaStreamNonBlocking));
        *out = c.own[c.device];
    }
    if (device) *device = c.device;
    return ok();
}

// Event pool.  A CUDA event belongs to the device that was current when it was created and can only be recorded on
// that device's streams, so the free lists are per device (one process may drive several GPUs through afk_set_device).
static std::mutex g_ev_mu;
static std::vector<cudaEvent_t> g_ev_pool[16];
static std::unordered_map<cudaEvent_t, int> g_ev_device;

cudaEvent_t event_get() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 16) {
        cudaGetLastError();
        return nullptr;
    }
    {
        std::lock_guard<std::mutex> lk(g_ev_mu);
        if (!g_ev_pool[dev].empty()) {
            cudaEvent_t e = g_ev_pool[dev].back();
            g_ev_pool[dev].pop_back();
            return e;
        }
    }
    cudaEvent_t e = nullptr;
    if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    std::lock_guard<std::mutex> lk(g_ev_mu);
    g_ev_device[e] = dev;
    return e;
}

void event_put(cudaEvent_t e) {
    if (!e) return;
    std::lock_guard<std::mutex> lk(g_ev_mu);
    auto it = g_ev_device.find(e);
    if (it == g_ev_device.end()) return; // not one of ours
    g_ev_pool[it->second].push_back(e);
}

// ---- deferred panics -------------------------------------------------------------------------------------------
Error panic_word_get(int device, unsigned **out) {
    ThreadCtx &c = tls();
    if (device < 0 || device >= 16) return make_error(AFK_ERR_CUDA, "CUDA: device index out of range");
    if (!c.panic_word[device]) {
        void *p = nullptr;
        AFK_CUDA(cudaHostAlloc(&p, 64, cudaHostAllocMapped | cudaHostAllocPortable));
        memset(p, 0, 64);
        c.panic_word[device] = static_cast<unsigned *>(p);
    }
    *out = c.panic_word[device];
    return ok();
}

Error panic_word_take(int device) {
    ThreadCtx &c = tls();
    if (device < 0 || device >= 16 || !c.panic_word[device]) return ok();
    volatile unsigned *w = c.panic_word[device];
    if (*w == 0u) return ok();
    *w = 0u;
    const size_t len = c.panic_tone_len;
    return make_error(AFK_ERR_PANIC, fmt("index out of bounds: the len is %zu but the index is larger (input above 65535*(1+1/%zu))", len,
                                         len > 0 ? len - 1 : 0));
}

// ---- read-back page --------------------------------------------------------------------------------------------
Error readback_page(char **out) {
    ThreadCtx &c = tls();
    if (!c.readback) {
        void *p = nullptr;
        AFK_CUDA(cudaHostAlloc(&p, kReadbackBytes, cudaHostAllocMapped | cudaHostAllocPortable));
        memset(p, 0, kReadbackBytes);
        c.readback = static_cast<char *>(p);
    }
    *out = c.readback;
    return ok();
}

// ---- scratch ---------------------------------------------------------------------------------------------
Error scratch_u32(cudaStream_t s, size_t n, unsigned **out) {
    void *p = nullptr;
    AFK_CUDA(cudaMallocAsync(&p, n * sizeof(unsigned), s));
    AFK_CUDA(cudaMemsetAsync(p, 0, n * sizeof(unsigned), s));
    *out = static_cast<unsigned *>(p);
    return ok();
}

void scratch_free(cudaStream_t s, void *p) {
    if (p) cudaFreeAsync(p, s);
}

Error reduce_workspace(const ReducePlan &plan, int device, cudaStream_t stream, ReduceWorkspace *out) {
    StreamScratch &sc = tls().scratch[std::make_pair(device, stream)];
    ReduceWorkspace &ws = sc.ws;
    if (ws.partial_bytes < plan.partial_bytes) {
        void *old = ws.partials; // forget the block before anything can fail: a failed allocation must not leave a freed pointer behind
        ws.partials = nullptr;
        ws.partial_bytes = 0;
        if (old) AFK_CUDA(cudaFreeAsync(old, stream));
        size_t want = plan.partial_bytes + plan.partial_bytes / 4;
        AFK_CUDA(cudaMallocAsync(&ws.partials, want, stream));
        ws.partial_bytes = want;
    }
    if (ws.n_counters < plan.n_counters) {
        void *old = ws.counters;
        ws.counters = nullptr;
        ws.n_counters = 0;
        if (old) AFK_CUDA(cudaFreeAsync(old, stream));
        size_t want = plan.n_counters * 2 + 64;
        void *p = nullptr;
        AFK_CUDA(cudaMallocAsync(&p, want * sizeof(unsigned), stream));
        AFK_CUDA(cudaMemsetAsync(p, 0, want * sizeof(unsigned), stream)); // counters reset themselves afterwards
        ws.counters = static_cast<unsigned *>(p);
        ws.n_counters = want;
    }
    *out = ws;
    return ok();
}

// ---- images ------------------------------------------------------------------------------------------------
bool ImageMeta::meta_equal(const ImageMeta &o) const {
    // PartialEq of Option<MetaWcsArray> (unit.rs:210-218)
    if (has_meta != o.has_meta) return false;
    if (!has_meta) return true;
    for (int i = 0; i < 4; ++i)
        if (!(axes[i] == o.axes[i])) return false;
    return wcs == o.wcs;
}

static Error check_shape(int ndim, const int64_t *shape) {
    if (ndim < 0 || ndim > AFK_MAX_NDIM) return make_error(AFK_ERR_SHAPE, fmt("unsupported number of dimensions %d", ndim));
    int64_t n = 1;
    for (int i = 0; i < ndim; ++i) {
        if (shape[i] < 0) return make_error(AFK_ERR_SHAPE, "negative dimension");
        // the element count sizes the allocation and the kernels' index arithmetic: it must not wrap (FITS headers are
        // untrusted input) and 2^40 elements (4 TiB of f32) is beyond any device
        if (__builtin_mul_overflow(n, shape[i], &n) || n > (int64_t(1) << 40)) return make_error(AFK_ERR_SHAPE, "image too large");
    }
    return ok();
}

Error image_new(int ndim, const int64_t *shape, afk_image **out) {
    if (Error e = check_shape(ndim, shape)) return e;
    cudaStream_t s;
    int dev;
    if (Error e = current_stream(&s, &dev)) return e;
    afk_image *img = new afk_image();
    img->ndim = ndim;
    for (int i = 0; i < ndim; ++i) img->shape[i] = shape[i];
    img->device = dev;
    img->stream = s;
    img->bytes = (size_t)img->len() * sizeof(float);
    void *p = nullptr;
    if (trace_enabled() && img->bytes >= ((size_t)64 << 20)) {
        cudaMemPool_t pool, cur;
        uint64_t reserved = 0, used = 0, thr = 0;
        cudaDeviceGetDefaultMemPool(&pool, dev);
        cudaDeviceGetMemPool(&cur, dev);
        if (g_big_pool[dev]) pool = g_big_pool[dev];
        cudaMemPoolGetAttribute(pool, cudaMemPoolAttrReservedMemCurrent, &reserved);
        cudaMemPoolGetAttribute(pool, cudaMemPoolAttrUsedMemCurrent, &used);
        cudaMemPoolGetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr);
        fprintf(stderr, "[afk] pool reserved %llu MiB used %llu MiB threshold %llx default==current %d\n",
                (unsigned long long)(reserved >> 20), (unsigned long long)(used >> 20), (unsigned long long)thr, pool == cur);
    }
    const auto t0 = std::chrono::steady_clock::now();
    const size_t data_bytes = (img->bytes + 15) / 16 * 16, alloc_bytes = data_bytes + 16; // + the range memo (host.h)
    cudaError_t ce = (img->bytes >= kBigAllocBytes && g_big_pool[dev]) ? cudaMallocFromPoolAsync(&p, alloc_bytes, g_big_pool[dev], s)
                                                                       : cudaMallocAsync(&p, alloc_bytes, s);
    if (trace_enabled() && img->bytes >= ((size_t)64 << 20))
        fprintf(stderr, "[afk] cudaMallocAsync(%zu MiB) took %.2f ms\n", img->bytes >> 20,
                std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
    if (ce != cudaSuccess) {
        delete img;
        return cuda_error(ce, "cudaMallocAsync(image)");
    }
    img->ptr = static_cast<float *>(p);
    img->range = reinterpret_cast<unsigned *>(static_cast<char *>(p) + data_bytes);
    img->ready = event_get();
    *out = img;
    return ok();
}

Error image_wait(const afk_image *img, cudaStream_t stream) {
    if (img->stream != stream && img->ready) AFK_CUDA(cudaStreamWaitEvent(stream, img->ready, 0));
    return ok();
}

void image_note_read(const afk_image *cimg, cudaStream_t stream) {
    afk_image *img = const_cast<afk_image *>(cimg);
    while (img->parent) img = img->parent; // the allocation belongs to the root
    if ((!img->owned && !img->track_readers) || img->stream == stream) return;
    cudaEvent_t e = event_get();
    if (!e) return;
    if (cudaEventRecord(e, stream) != cudaSuccess) {
        cudaGetLastError();
        event_put(e);
        return;
    }
    std::lock_guard<std::mutex> lk(img->mu);
    img->readers.push_back(e);
    if (img->readers.size() > 64) {
        // bound the list: the owner stream takes over the dependency on the oldest read before its event is recycled,
        // so the stream-ordered free in image_destroy still comes after every reader (reads may come from many streams)
        cudaEvent_t old = img->readers.front();
        if (cudaEventQuery(old) != cudaSuccess) {
            cudaGetLastError();
            int prev = -1;
            cudaGetDevice(&prev);
            if (prev != img->device) cudaSetDevice(img->device);
            cudaStreamWaitEvent(img->stream, old, 0);
            if (prev >= 0 && prev != img->device) cudaSetDevice(prev);
        }
        event_put(old);
        img->readers.erase(img->readers.begin());
    }
}

Error image_mark_ready(afk_image *img, cudaStream_t stream) {
    img->stream = stream;
    if (!img->ready) img->ready = event_get();
    if (img->ready) AFK_CUDA(cudaEventRecord(img->ready, stream));
    return ok();
}

bool trace_enabled() {
    static const bool on = getenv("AFK_TRACE") != nullptr;
    return on;
}

static void image_destroy(afk_image *img) {
    if (img->parent) {
        afk_image_release(img->parent);
    } else if (img->ipc_peer) {
        cudaIpcCloseMemHandle(img->ptr);
    } else if (img->owned && img->ptr) {
        int prev = -1;
        cudaGetDevice(&prev);
        cudaSetDevice(img->device);
        for (cudaEvent_t e : img->readers) cudaStreamWaitEvent(img->stream, e, 0);
        const auto t0 = std::chrono::steady_clock::now();
        if (cudaFreeAsync(img->ptr, img->stream) != cudaSuccess) {
            cudaGetLastError();
            cudaFree(img->ptr);
        }
        if (trace_enabled() && img->bytes >= ((size_t)64 << 20))
            fprintf(stderr, "[afk] cudaFreeAsync(%zu MiB) took %.2f ms\n", img->bytes >> 20,
                    std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
        if (prev >= 0 && prev != img->device) cudaSetDevice(prev);
    }
    for (cudaEvent_t e : img->readers) event_put(e);
    event_put(img->ready);
    delete img;
}

} // namespace afk

using namespace afk;

// =============================================================================================================
// C ABI
// =============================================================================================================
extern "C" {

const char *afk_last_error(void) { return tls().last_error.c_str(); }
int afk_abi_version(void) { return AFK_ABI_VERSION; }

int afk_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

int afk_set_device(int device) {
    int n = afk_device_count();
    if (device < 0 || device >= n || device >= 16) return fail(make_error(AFK_ERR_CUDA, fmt("CUDA: invalid device %d (have %d)", device, n)));
    tls().device = device;
    cudaStream_t s;
    return fail(current_stream(&s, nullptr));
}

int afk_set_stream(void *cuda_stream) {
    tls().external = static_cast<cudaStream_t>(cuda_stream);
    return AFK_OK;
}

void *afk_get_stream(void) {
    cudaStream_t s = nullptr;
    if (fail(current_stream(&s, nullptr)) != AFK_OK) return nullptr;
    return s;
}

int afk_synchronize(void) {
    cudaStream_t s;
    if (Error e = current_stream(&s, nullptr)) return fail(e);
    cudaError_t ce = cudaStreamSynchronize(s);
    if (ce != cudaSuccess) return fail(cuda_error(ce, "cudaStreamSynchronize"));
    // a display chain enqueued by afk_colormap that ran into the reference's tone-table panic reports it here
    if (Error e = panic_word_take(tls().device)) return fail(e);
    return AFK_OK;
}

uint64_t afk_launch_count(void) { return launch_count(); }

// Sum mode: a process-wide default plus a per-thread override.  cake calls operators from arbitrary rayon workers
// (aflak_cake/src/dst/compute.rs:206-224), so a host that wants the reference's bits sets the default once.
int afk_set_default_sum_mode(int mode) {
    if (mode != SUM_COMPENSATED && mode != SUM_SEQUENTIAL) return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "unknown sum mode"));
    g_default_sum_mode.store(mode, std::memory_order_relaxed);
    return AFK_OK;
}
int afk_set_sum_mode(int mode) {
    if (mode != SUM_INHERIT && mode != SUM_COMPENSATED && mode != SUM_SEQUENTIAL) return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "unknown sum mode"));
    tls().sum_mode = mode;
    return AFK_OK;
}
int afk_get_sum_mode(void) { return effective_sum_mode(); }

// Return every cached (freed) block of the stream-ordered pools of the current device to the driver.  A host calls this
// before an allocation that needs nearly the whole device (the 128 GiB cube of BASELINE configs[4] on one GPU).
int afk_trim_pools(void) {
    cudaStream_t s;
    int dev;
    if (Error e = current_stream(&s, &dev)) return fail(e);
    cudaError_t ce = cudaDeviceSynchronize();
    cudaMemPool_t pool;
    if (ce == cudaSuccess) ce = cudaDeviceGetDefaultMemPool(&pool, dev);
    if (ce == cudaSuccess) ce = cudaMemPoolTrimTo(pool, 0);
    if (ce == cudaSuccess && g_big_pool[dev]) ce = cudaMemPoolTrimTo(g_big_pool[dev], 0);
    if (ce != cudaSuccess) return fail(cuda_error(ce, "afk_trim_pools"));
    return AFK_OK;
}

// ---- type ids ---------------------------------------------------------------------------------------------
static const char *kTypeNames[AFK_T_COUNT] = {"Integer", "Float", "Float2", "Float3", "Float3x3", "Str", "Bool",
                                              "Fits", "Image", "Map2dTo3dCoords", "Roi", "Paths", "ToneCurve"};

const char *afk_type_name(afk_type t) { return (t >= 0 && t < AFK_T_COUNT) ? kTypeNames[t] : ""; }

int afk_type_from_name(const char *name) {
    if (!name) return -1;
    for (int i = 0; i < AFK_T_COUNT; ++i)
        if (strcmp(kTypeNames[i], name) == 0) return i;
    return -1;
}

int afk_type_convertible(afk_type from, afk_type into) {
    if (from == into) return 1;
    return (from == AFK_T_INTEGER && into == AFK_T_FLOAT) || (from == AFK_T_FLOAT && into == AFK_T_INTEGER);
}

int afk_type_editable(afk_type t) {
    switch (t) {
    case AFK_T_INTEGER: case AFK_T_FLOAT: case AFK_T_FLOAT2: case AFK_T_FLOAT3: case AFK_T_FLOAT3X3:
    case AFK_T_ROI: case AFK_T_STR: case AFK_T_BOOL: case AFK_T_PATHS: case AFK_T_TONECURVE:
        return 1;
    default:
        return 0;
    }
}

// ---- images -------------------------------------------------------------------------------------------------
afk_image *afk_image_alloc(int ndim, const int64_t *shape) {
    afk_image *img = nullptr;
    if (fail(image_new(ndim, shape, &img)) != AFK_OK) return nullptr;
    cudaStream_t s;
    if (fail(current_stream(&s, nullptr)) != AFK_OK || fail(image_mark_ready(img, s)) != AFK_OK) {
        afk_image_release(img);
        return nullptr;
    }
    return img;
}

afk_image *afk_image_upload(const float *host, int ndim, const int64_t *shape) {
    if (!host) {
        fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "null host pointer"));
        return nullptr;
    }
    afk_image *img = nullptr;
    if (fail(image_new(ndim, shape, &img)) != AFK_OK) return nullptr;
    cudaStream_t s = img->stream;
    // pinned-memory cudaMemcpyAsync (north_star subsystem 2): asynchronous when `host` is page-locked, staged by the
    // driver otherwise.  Large cubes are copied in frame-aligned chunks so the copy engine pipelines them.
    const size_t chunk = (size_t)256 << 20;
    const char *src = reinterpret_cast<const char *>(host);
    char *dst = reinterpret_cast<char *>(img->ptr);
    const auto t0 = std::chrono::steady_clock::now();
    for (size_t off = 0; off < img->bytes; off += chunk) {
        size_t n = img->bytes - off < chunk ? img->bytes - off : chunk;
        cudaError_t ce = cudaMemcpyAsync(dst + off, src + off, n, cudaMemcpyHostToDevice, s);
        if (ce != cudaSuccess) {
            fail(cuda_error(ce, "cudaMemcpyAsync(H2D)"));
            afk_image_release(img);
            return nullptr;
        }
    }
    if (trace_enabled() && img->bytes >= ((size_t)64 << 20))
        fprintf(stderr, "[afk] H2D enqueue of %zu MiB took %.2f ms\n", img->bytes >> 20,
                std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
    if (fail(image_mark_ready(img, s)) != AFK_OK) {
        afk_image_release(img);
        return nullptr;
    }
    return img;
}

afk_image *afk_image_wrap_device(float *device_ptr, int ndim, const int64_t *shape) {
    if (ndim < 0 || ndim > AFK_MAX_NDIM || !device_ptr || (ndim > 0 && !shape)) {
        fail(make_error(AFK_ERR_SHAPE, "bad arguments to afk_image_wrap_device"));
        return nullptr;
    }
    if (fail(check_shape(ndim, shape)) != AFK_OK) return nullptr; // negative or overflowing extents: the kernels index with the product
    cudaStream_t s;
    int dev;
    if (fail(current_stream(&s, &dev)) != AFK_OK) return nullptr;
    afk_image *img = new afk_image();
    img->ndim = ndim;
    for (int i = 0; i < ndim; ++i) img->shape[i] = shape[i];
    img->ptr = device_ptr;
    img->owned = false;
    img->device = dev;
    img->stream = s; // assumed produced on (or already synchronised with) the caller's current stream
    img->bytes = (size_t)img->len() * sizeof(float);
    return img;
}

afk_image *afk_image_synthetic(uint64_t seed, int64_t nz, int64_t ny, int64_t nx, int64_t y0, int64_t rows, int pathology) {
    int64_t shape[3] = {nz, rows, nx};
    afk_image *img = nullptr;
    if (fail(image_new(3, shape, &img)) != AFK_OK) return nullptr;
    cudaError_t ce = gen_cube_launch(seed, nz, ny, nx, y0, rows, pathology, img->ptr, img->stream);
    if (ce != cudaSuccess) {
        fail(cuda_error(ce, "gen_cube"));
        afk_image_release(img);
        return nullptr;
    }
    if (fail(image_mark_ready(img, img->stream)) != AFK_OK) {
        afk_image_release(img);
        return nullptr;
    }
    return img;
}

void afk_image_retain(afk_image *img) {
    if (img) img->rc.fetch_add(1, std::memory_order_relaxed);
}

void afk_image_release(afk_image *img) {
    if (img && img->rc.fetch_sub(1, std::memory_order_acq_rel) == 1) image_destroy(img);
}

int afk_image_ndim(const afk_image *img) { return img ? img->ndim : 0; }
const int64_t *afk_image_shape(const afk_image *img) { return img ? img->shape : nullptr; }
int64_t afk_image_len(const afk_image *img) { return img ? img->len() : 0; }
float *afk_image_device_ptr(const afk_image *img) { return img ? img->ptr : nullptr; }
int afk_image_device(const afk_image *img) { return img ? img->device : -1; }

int afk_image_download(const afk_image *img, float *host) {
    if (!img || !host) return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "null argument to afk_image_download"));
    cudaStream_t s;
    if (Error e = current_stream(&s, nullptr)) return fail(e);
    if (Error e = image_wait(img, s)) return fail(e);
    if (img->bytes) {
        cudaError_t ce = cudaMemcpyAsync(host, img->ptr, img->bytes, cudaMemcpyDeviceToHost, s);
        if (ce != cudaSuccess) return fail(cuda_error(ce, "cudaMemcpyAsync(D2H)"));
    }
    cudaError_t ce = cudaStreamSynchronize(s);
    if (ce != cudaSuccess) return fail(cuda_error(ce, "cudaStreamSynchronize"));
    return AFK_OK;
}

const char *afk_image_unit(const afk_image *img) { return (img && img->meta.has_unit) ? img->meta.unit.c_str() : ""; }

int afk_image_set_unit(afk_image *img, const char *unit) {
    if (!img) return AFK_ERR_UNEXPECTED_INPUT;
    img->meta.has_unit = unit && unit[0];
    img->meta.unit = unit ? unit : "";
    return AFK_OK;
}

int afk_image_homogeneous(const afk_image *img) { return img ? (img->meta.homogeneous ? 1 : 0) : 0; }
const char *afk_image_tag(const afk_image *img) { return (img && img->meta.has_tag) ? img->meta.tag.c_str() : nullptr; }

int afk_image_set_tag(afk_image *img, const char *tag) {
    if (!img) return AFK_ERR_UNEXPECTED_INPUT;
    img->meta.has_tag = tag != nullptr;
    img->meta.tag = tag ? tag : "";
    return AFK_OK;
}

int afk_image_has_meta(const afk_image *img) { return img && img->meta.has_meta; }

int afk_image_set_axis(afk_image *img, int axis, const char *name, const char *unit) {
    if (!img || axis < 0 || axis >= 4) return AFK_ERR_UNEXPECTED_INPUT;
    img->meta.has_meta = true;
    img->meta.axes[axis].has_name = name != nullptr;
    img->meta.axes[axis].name = name ? name : "";
    img->meta.axes[axis].unit = unit ? unit : "";
    return AFK_OK;
}

int afk_image_set_wcs_axis(afk_image *img, int axis, double crval, double crpix, double cdelt) {
    if (!img || axis < 0 || axis >= 4) return AFK_ERR_UNEXPECTED_INPUT;
    img->meta.has_meta = true;
    if (img->meta.wcs.size() != 12) {
        img->meta.wcs.assign(12, 0.0);
        for (int i = 0; i < 4; ++i) img->meta.wcs[3 * i + 1] = img->meta.wcs[3 * i + 2] = 1.0; // identity
    }
    img->meta.wcs[3 * axis + 0] = crval;
    img->meta.wcs[3 * axis + 1] = crpix;
    img->meta.wcs[3 * axis + 2] = cdelt;
    return AFK_OK;
}

int afk_image_pix2world(const afk_image *img, int axis, float pixel, float *world) {
    if (!img || !world || axis < 0 || axis >= 4) return 0;
    if (!img->meta.has_meta) return 0; // Option::None (unit.rs:336)
    *world = afk::image_pix2world(img, axis, pixel);
    return 1;
}

const char *afk_image_axis_name(const afk_image *img, int axis) {
    return (img && img->meta.has_meta && axis >= 0 && axis < 4) ? img->meta.axes[axis].name.c_str() : "";
}
const char *afk_image_axis_unit(const afk_image *img, int axis) {
    return (img && img->meta.has_meta && axis >= 0 && axis < 4) ? img->meta.axes[axis].unit.c_str() : "";
}

// ---- values -----------------------------------------------------------------------------------------------------
static afk_value *new_value(afk_type t) {
    afk_value *v = new afk_value();
    v->type = t;
    return v;
}

afk_value *afk_value_integer(int64_t x) {
    afk_value *v = new_value(AFK_T_INTEGER);
    v->i = x;
    return v;
}
afk_value *afk_value_float(float x) {
    afk_value *v = new_value(AFK_T_FLOAT);
    v->f[0] = x;
    return v;
}
afk_value *afk_value_float2(const float x[2]) {
    afk_value *v = new_value(AFK_T_FLOAT2);
    memcpy(v->f, x, 2 * sizeof(float));
    return v;
}
afk_value *afk_value_float3(const float x[3]) {
    afk_value *v = new_value(AFK_T_FLOAT3);
    memcpy(v->f, x, 3 * sizeof(float));
    return v;
}
afk_value *afk_value_float3x3(const float x[9]) {
    afk_value *v = new_value(AFK_T_FLOAT3X3);
    memcpy(v->f, x, 9 * sizeof(float));
    return v;
}
afk_value *afk_value_str(const char *s) {
    afk_value *v = new_value(AFK_T_STR);
    v->s = s ? s : "";
    return v;
}
afk_value *afk_value_bool(int b) {
    afk_value *v = new_value(AFK_T_BOOL);
    v->b = b != 0;
    return v;
}
afk_value *afk_value_image(afk_image *img) {
    if (!img) return nullptr;
    afk_value *v = new_value(AFK_T_IMAGE);
    afk_image_retain(img);
    v->img = img;
    return v;
}
afk_value *afk_value_tone_curve(const float *arr, size_t n, const float *cp_xy, size_t n_cp) {
    afk_value *v = new_value(AFK_T_TONECURVE);
    if (arr) v->arr.assign(arr, arr + n);
    if (cp_xy) v->cp.assign(cp_xy, cp_xy + 2 * n_cp);
    return v;
}

afk_value *afk_value_roi_all(void) { return new_value(AFK_T_ROI); }
afk_value *afk_value_roi_pixels(const uint64_t *xy, size_t n_pixels) {
    afk_value *v = new_value(AFK_T_ROI);
    v->roi_all = false;
    if (xy) v->roi_xy.assign(xy, xy + 2 * n_pixels);
    return v;
}
int afk_value_roi_is_all(const afk_value *v) { return v && v->type == AFK_T_ROI && v->roi_all; }
size_t afk_value_roi_count(const afk_value *v) { return (v && v->type == AFK_T_ROI && !v->roi_all) ? v->roi_xy.size() / 2 : 0; }
const uint64_t *afk_value_roi_data(const afk_value *v) { return (v && v->type == AFK_T_ROI && !v->roi_all) ? v->roi_xy.data() : nullptr; }
afk_value *afk_value_map2dto3d(const float *xyz, int64_t rows, int64_t cols) {
    if (rows < 0 || cols < 0 || (!xyz && rows * cols > 0)) return nullptr;
    afk_value *v = new_value(AFK_T_MAP2DTO3DCOORDS);
    v->map_rows = rows;
    v->map_cols = cols;
    if (xyz) v->map.assign(xyz, xyz + 3 * rows * cols);
    return v;
}
const float *afk_value_get_map(const afk_value *v, int64_t *rows, int64_t *cols) {
    if (!v || v->type != AFK_T_MAP2DTO3DCOORDS) return nullptr;
    if (rows) *rows = v->map_rows;
    if (cols) *cols = v->map_cols;
    return v->map.data();
}

afk_value *afk_value_default_for(afk_type t) {
    // DefaultFor (lib.rs:760-776)
    switch (t) {
    case AFK_T_INTEGER: return afk_value_integer(0);
    case AFK_T_FLOAT: return afk_value_float(0.0f);
    case AFK_T_FLOAT2: { float z[2] = {0, 0}; return afk_value_float2(z); }
    case AFK_T_FLOAT3: { float z[3] = {0, 0, 0}; return afk_value_float3(z); }
    case AFK_T_FLOAT3X3: { float z[9] = {0}; return afk_value_float3x3(z); }
    case AFK_T_STR: return afk_value_str("");
    case AFK_T_BOOL: return afk_value_bool(0);
    case AFK_T_ROI: return new_value(AFK_T_ROI);     // ROI::All
    case AFK_T_PATHS: return new_value(AFK_T_PATHS); // FileList(vec![])
    case AFK_T_TONECURVE: {
        // ToneCurveState::default (imgui_tone_curve/src/lib.rs:21-38): arr[i] = i, cp = (0,0),(1,1)
        float arr[256];
        for (int i = 0; i < 256; ++i) arr[i] = (float)i;
        float cp[4] = {0.0f, 0.0f, 1.0f, 1.0f};
        return afk_value_tone_curve(arr, 256, cp, 2);
    }
    default:
        fail(make_error(AFK_ERR_PANIC, fmt("Unknown variant name provided: %s.", afk_type_name(t))));
        return nullptr;
    }
}

void afk_value_retain(afk_value *v) {
    if (v) v->rc.fetch_add(1, std::memory_order_relaxed);
}

void afk_value_release(afk_value *v) {
    if (v && v->rc.fetch_sub(1, std::memory_order_acq_rel) == 1) {
        if (v->img) afk_image_release(v->img);
        if (v->fits) afk_fits_release(v->fits);
        if (v->map_dev) cudaFree(v->map_dev);
        if (v->roi_dev) cudaFree(v->roi_dev);
        delete v;
    }
}

afk_type afk_value_type(const afk_value *v) { return v ? v->type : AFK_T_COUNT; }

int afk_value_get_integer(const afk_value *v, int64_t *out) {
    if (!v || v->type != AFK_T_INTEGER) return AFK_ERR_PANIC;
    *out = v->i;
    return AFK_OK;
}
int afk_value_get_float(const afk_value *v, float *out) {
    if (!v || v->type != AFK_T_FLOAT) return AFK_ERR_PANIC;
    *out = v->f[0];
    return AFK_OK;
}
int afk_value_get_floats(const afk_value *v, float *out, size_t cap) {
    if (!v) return 0;
    const float *src = v->f;
    size_t n = 0;
    switch (v->type) {
    case AFK_T_FLOAT: n = 1; break;
    case AFK_T_FLOAT2: n = 2; break;
    case AFK_T_FLOAT3: n = 3; break;
    case AFK_T_FLOAT3X3: n = 9; break;
    case AFK_T_TONECURVE: n = v->arr.size(); src = v->arr.data(); break;
    default: return 0;
    }
    for (size_t i = 0; i < n && i < cap; ++i) out[i] = src[i];
    return (int)n;
}
int afk_value_get_bool(const afk_value *v, int *out) {
    if (!v || v->type != AFK_T_BOOL) return AFK_ERR_PANIC;
    *out = v->b ? 1 : 0;
    return AFK_OK;
}
const char *afk_value_get_str(const afk_value *v) { return (v && v->type == AFK_T_STR) ? v->s.c_str() : nullptr; }
afk_image *afk_value_get_image(const afk_value *v) { return (v && v->type == AFK_T_IMAGE) ? v->img : nullptr; }

afk_value *afk_value_convert(const afk_value *v, afk_type into) {
    if (!v) return nullptr;
    if (v->type == into) {
        afk_value_retain(const_cast<afk_value *>(v)); // Bow::Borrowed
        return const_cast<afk_value *>(v);
    }
    if (v->type == AFK_T_INTEGER && into == AFK_T_FLOAT) return afk_value_float((float)v->i); // integer_to_float, lib.rs:810-816
    if (v->type == AFK_T_FLOAT && into == AFK_T_INTEGER) {
        // float_to_integer, lib.rs:817-823: f.round() as i64 (round half away from zero, saturating, NaN -> 0)
        float r = roundf(v->f[0]);
        int64_t i;
        if (r != r) i = 0;
        else if (r >= 9223372036854775808.0f) i = INT64_MAX;
        else if (r <= -9223372036854775808.0f) i = INT64_MIN;
        else i = (int64_t)r;
        return afk_value_integer(i);
    }
    return nullptr;
}

static bool image_eq(const afk_image *a, const afk_image *b) {
    if (a == b) {
        // same handle: equal unless it contains a NaN (f32 PartialEq); checked below like any other pair
    }
    if (a->ndim != b->ndim) return false;
    for (int i = 0; i < a->ndim; ++i)
        if (a->shape[i] != b->shape[i]) return false;
    const ImageMeta &ma = a->meta, &mb = b->meta;
    if (ma.has_unit != mb.has_unit || (ma.has_unit && ma.unit != mb.unit) || ma.homogeneous != mb.homogeneous) return false;
    if (ma.has_tag != mb.has_tag || (ma.has_tag && ma.tag != mb.tag)) return false;
    if (!ma.meta_equal(mb)) return false;
    size_t n = (size_t)a->len();
    std::vector<float> ha(n), hb(n);
    if (afk_image_download(a, ha.data()) != AFK_OK || afk_image_download(b, hb.data()) != AFK_OK) return false;
    for (size_t i = 0; i < n; ++i)
        if (!(ha[i] == hb[i])) return false;
    return true;
}

int afk_value_eq(const afk_value *a, const afk_value *b) {
    // PartialEq for IOValue (lib.rs:85-104); Fits never compares equal
    if (!a || !b || a->type != b->type) return 0;
    switch (a->type) {
    case AFK_T_INTEGER: return a->i == b->i;
    case AFK_T_FLOAT: return a->f[0] == b->f[0];
    case AFK_T_FLOAT2: return a->f[0] == b->f[0] && a->f[1] == b->f[1];
    case AFK_T_FLOAT3: return a->f[0] == b->f[0] && a->f[1] == b->f[1] && a->f[2] == b->f[2];
    case AFK_T_FLOAT3X3:
        for (int i = 0; i < 9; ++i)
            if (!(a->f[i] == b->f[i])) return 0;
        return 1;
    case AFK_T_STR: return a->s == b->s;
    case AFK_T_BOOL: return a->b == b->b;
    case AFK_T_IMAGE: return image_eq(a->img, b->img);
    case AFK_T_TONECURVE: return a->cp == b->cp; // control points only (imgui_tone_curve/src/lib.rs:50-56)
    case AFK_T_PATHS: return a->paths == b->paths;
    case AFK_T_ROI: return a->roi_all == b->roi_all && (a->roi_all || a->roi_xy == b->roi_xy);
    case AFK_T_MAP2DTO3DCOORDS: return a->map_rows == b->map_rows && a->map_cols == b->map_cols && a->map == b->map;
    default: return 0;
    }
}

// ---- results --------------------------------------------------------------------------------------------------------
void afk_result_clear(afk_result *r) {
    if (!r) return;
    if (r->value) afk_value_release(r->value);
    free(r->message);
    r->value = nullptr;
    r->message = nullptr;
    r->status = AFK_OK;
}

void afk_free(void *p) { free(p); }

char *afk_result_display(const afk_result *r) {
    // Display for IOErr (lib.rs:118-130)
    if (!r) return nullptr;
    const char *m = r->message ? r->message : "";
    std::string s;
    switch (r->status) {
    case AFK_OK: s = "Ok"; break;
    case AFK_ERR_IO: s = fmt("I/O error! %s.", m); break;
    case AFK_ERR_RAWLOADER: s = fmt("RawLoader-related error! %s", m); break;
    case AFK_ERR_FITS: s = fmt("FITS-related error! %s", m); break;
    case AFK_ERR_UNEXPECTED_INPUT: case AFK_ERR_CUDA: s = fmt("Unexpected input! %s", m); break;
    case AFK_ERR_SHAPE: s = fmt("Shape error! %s.", m); break;
    case AFK_ERR_PANIC: s = fmt("panicked: %s", m); break;
    case AFK_ERR_ARGUMENT: s = m; break;
    }
    return strdup(s.c_str());
}

// ---- profiling of the reduce kernel (bench.py roofline) -----------------------------------------------------------
int afk_profile_reduce(int enable) {
    ThreadCtx &c = tls();
    for (auto &p : c.prof_events) {
        cudaEventDestroy(p.first);
        cudaEventDestroy(p.second);
    }
    c.prof_events.clear();
    c.prof_launches = 0;
    c.profile = enable != 0;
    return AFK_OK;
}

int afk_profile_reduce_read(double *total_ms, uint64_t *launches) {
    ThreadCtx &c = tls();
    double tot = 0.0;
    for (auto &p : c.prof_events) {
        if (cudaEventSynchronize(p.second) != cudaSuccess) return fail(cuda_error(cudaGetLastError(), "cudaEventSynchronize"));
        float ms = 0.0f;
        if (cudaEventElapsedTime(&ms, p.first, p.second) != cudaSuccess) return fail(cuda_error(cudaGetLastError(), "cudaEventElapsedTime"));
        tot += ms;
    }
    if (total_ms) *total_ms = tot;
    if (launches) *launches = c.prof_launches;
    return AFK_OK;
}

// ---- CUDA IPC (one process per GPU; fused gather writes bands straight into the root's buffer) --------------------
int afk_ipc_export(const afk_image *img, uint8_t handle[64]) {
    if (!img || !handle) return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "null argument to afk_ipc_export"));
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
    cudaIpcMemHandle_t h;
    cudaError_t ce = cudaIpcGetMemHandle(&h, img->ptr);
    if (ce != cudaSuccess) return fail(cuda_error(ce, "cudaIpcGetMemHandle"));
    memcpy(handle, &h, 64);
    return AFK_OK;
}

float *afk_ipc_open(const uint8_t handle[64]) {
    cudaStream_t s;
    if (fail(current_stream(&s, nullptr)) != AFK_OK) return nullptr;
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, 64);
    void *p = nullptr;
    cudaError_t ce = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
    if (ce != cudaSuccess) {
        fail(cuda_error(ce, "cudaIpcOpenMemHandle"));
        return nullptr;
    }
    return static_cast<float *>(p);
}

int afk_ipc_close(float *peer_ptr) {
    if (!peer_ptr) return AFK_OK;
    cudaError_t ce = cudaIpcCloseMemHandle(peer_ptr);
    if (ce != cudaSuccess) return fail(cuda_error(ce, "cudaIpcCloseMemHandle"));
    return AFK_OK;
}

} // extern "C"
