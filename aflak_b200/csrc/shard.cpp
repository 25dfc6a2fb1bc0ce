// shard.cpp -- spatial sharding of a cube across the GPUs of one box, and the fused gather of the 2-D result.
//
// The reference has no distribution at all (single process, rayon).  Every output pixel of the hot path
// depends only on its own spaxel, so a cube [nz, ny, nx] shards into contiguous y-bands, one per GPU, with no
// data-path collective; the only exchange is the gather of the [ny, nx] result.  That gather is fused
// into the reduction kernel: each rank's kernel stores its result band straight into every rank's copy of the full
// image through CUDA-IPC peer mappings (NVLink 5 / NVSwitch) and the last CTA publishes an epoch flag; consumers wait
// on the flags with a one-warp kernel in stream order.  No NCCL on the data path; the host side's process group
// (torch.distributed) only carries the 64-byte IPC handles once at start-up.
#include <cstring>

#include "host.h"

struct afk_gather {
    int world = 1, rank = 0, device = 0;
    int64_t ny = 0, nx = 0;
    int64_t y0[afk::kMaxPeers] = {0}, rows[afk::kMaxPeers] = {0};
    size_t slot_floats = 0;         // one full image, padded to 256 B
    // cudaMalloc'ed: [slot 0][slot 1][rgba slot 0][rgba slot 1][control words].  Control words (u32 index):
    //   0..7 flags | 8 done | 9 err | 16..23 rgba flags | 24 rgba done | 25 rgba err
    //   32.. range table [2 slots][kMaxPeers][4] (min/max exchange) | 96.. combined range [2 slots][4]
    char *local = nullptr;
    char *peer[afk::kMaxPeers] = {nullptr};
    char *mc = nullptr;             // NVLink-SHARP multicast mapping of the same layout (symmetric memory from the host), or nullptr
    bool owned = true;              // false: the memory belongs to the host (afk_gather_create_symmetric)
    bool connected = false;
    unsigned epoch = 0;        // last epoch a reduce was enqueued for
    unsigned waited_epoch = 0; // last epoch a wait was enqueued for
    cudaStream_t stream = nullptr; // every reduce / wait of this gather is ordered on this stream
    afk_image *result[2] = {nullptr, nullptr}; // borrowed-memory handles of the two slots (readers on other streams are tracked)
    size_t bytes = 0;
    unsigned rgba_epoch = 0, rgba_waited = 0; // RGBA band gather of the sharded display chain (own flags, own two slots)
    static constexpr size_t kControlWords = 128;
    float *slot(int r, unsigned e) const { return reinterpret_cast<float *>(peer[r]) + (size_t)(e & 1u) * slot_floats; }
    uint8_t *rgba_slot(int r, unsigned e) const {
        return reinterpret_cast<uint8_t *>(peer[r]) + (2 + (size_t)(e & 1u)) * slot_floats * sizeof(float);
    }
    unsigned *control(int r) const { return reinterpret_cast<unsigned *>(peer[r] + 4 * slot_floats * sizeof(float)); }
    unsigned *flags(int r) const { return control(r); }
    unsigned *done() const { return control(rank) + 8; }
    unsigned *err() const { return control(rank) + 9; }
    unsigned *rgba_flags(int r) const { return control(r) + 16; }
    unsigned *rgba_done() const { return control(rank) + 24; }
    unsigned *rgba_err() const { return control(rank) + 25; }
    unsigned *range_entry(int r, unsigned e, int of_rank) const { return control(r) + 32 + ((size_t)(e & 1u) * afk::kMaxPeers + of_rank) * 4; }
    unsigned *combined(unsigned e) const { return control(rank) + 96 + (size_t)(e & 1u) * 4; }
    float *mc_slot(unsigned e) const { return reinterpret_cast<float *>(mc) + (size_t)(e & 1u) * slot_floats; }
    uint8_t *mc_rgba_slot(unsigned e) const { return reinterpret_cast<uint8_t *>(mc) + (2 + (size_t)(e & 1u)) * slot_floats * sizeof(float); }
};

static size_t gather_slot_floats(int64_t ny_total, int64_t nx) { return ((size_t)ny_total * (size_t)nx + 63) / 64 * 64; }
static size_t gather_bytes(int64_t ny_total, int64_t nx) {
    return 4 * gather_slot_floats(ny_total, nx) * sizeof(float) + afk_gather::kControlWords * sizeof(unsigned);
}

using namespace afk;

extern "C" {

// Band of rank `rank`: rows split as evenly as possible, the first (ny % world) ranks get one more row.
int afk_shard_band(int64_t ny, int world, int rank, int64_t *y0, int64_t *rows) {
    if (world < 1 || rank < 0 || rank >= world || ny < 0 || !y0 || !rows) return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "bad arguments to afk_shard_band"));
    const int64_t base = ny / world, rem = ny % world;
    *y0 = rank * base + (rank < rem ? rank : rem);
    *rows = base + (rank < rem ? 1 : 0);
    return AFK_OK;
}

afk_gather *afk_gather_create(int world, int rank, int64_t ny_total, int64_t nx, uint8_t handle_out[64]) {
    if (world < 1 || world > kMaxPeers || rank < 0 || rank >= world || ny_total < 0 || nx < 0 || !handle_out) {
        fail(make_error(AFK_ERR_UNEXPECTED_INPUT, fmt("afk_gather_create: world must be 1..%d and 0 <= rank < world", kMaxPeers)));
        return nullptr;
    }
    cudaStream_t s;
    int dev;
    if (fail(current_stream(&s, &dev)) != AFK_OK) return nullptr;
    afk_gather *g = new afk_gather();
    g->world = world;
    g->rank = rank;
    g->device = dev;
    g->stream = s;
    g->ny = ny_total;
    g->nx = nx;
    for (int r = 0; r < world; ++r) afk_shard_band(ny_total, world, r, &g->y0[r], &g->rows[r]);
    g->slot_floats = gather_slot_floats(ny_total, nx);
    g->bytes = gather_bytes(ny_total, nx);
    void *p = nullptr;
    // cudaMalloc, not the stream-ordered pool: pool memory cannot be exported with cudaIpcGetMemHandle
    cudaError_t ce = cudaMalloc(&p, g->bytes);
    if (ce == cudaSuccess) ce = cudaMemset(p, 0, g->bytes);
    cudaIpcMemHandle_t h;
    if (ce == cudaSuccess) ce = cudaIpcGetMemHandle(&h, p);
    if (ce != cudaSuccess) {
        fail(cuda_error(ce, "afk_gather_create"));
        if (p) cudaFree(p);
        delete g;
        return nullptr;
    }
    g->local = static_cast<char *>(p);
    g->peer[rank] = g->local;
    memcpy(handle_out, &h, 64);
    if (world == 1) g->connected = true;
    return g;
}

size_t afk_gather_bytes(int64_t ny_total, int64_t nx) {
    if (ny_total < 0 || nx < 0) return 0;
    return gather_bytes(ny_total, nx);
}

// The same gather over memory the HOST supplies: a symmetric allocation of afk_gather_bytes() zeroed bytes on every rank,
// peer_bases[r] = rank r's buffer as mapped in this process (peer_bases[rank] = the local one), multicast_base = an
// NVLink-SHARP multicast mapping of the same buffers or NULL.  With a multicast mapping the reduce kernel's epilogue emits ONE
// multimem.st per tile instead of one store per GPU (the switch replicates it), and so does the RGBA band broadcast.
afk_gather *afk_gather_create_symmetric(int world, int rank, int64_t ny_total, int64_t nx, void *const *peer_bases, void *multicast_base) {
    if (world < 1 || world > kMaxPeers || rank < 0 || rank >= world || ny_total < 0 || nx < 0 || !peer_bases) {
        fail(make_error(AFK_ERR_UNEXPECTED_INPUT, fmt("afk_gather_create_symmetric: world must be 1..%d and 0 <= rank < world", kMaxPeers)));
        return nullptr;
    }
    for (int r = 0; r < world; ++r)
        if (!peer_bases[r] || (reinterpret_cast<uintptr_t>(peer_bases[r]) & 255u)) {
            fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "afk_gather_create_symmetric: every peer base must be a non-null, 256-byte aligned device pointer"));
            return nullptr;
        }
    cudaStream_t s;
    int dev;
    if (fail(current_stream(&s, &dev)) != AFK_OK) return nullptr;
    afk_gather *g = new afk_gather();
    g->world = world;
    g->rank = rank;
    g->device = dev;
    g->stream = s;
    g->ny = ny_total;
    g->nx = nx;
    for (int r = 0; r < world; ++r) afk_shard_band(ny_total, world, r, &g->y0[r], &g->rows[r]);
    g->slot_floats = gather_slot_floats(ny_total, nx);
    g->bytes = gather_bytes(ny_total, nx);
    for (int r = 0; r < world; ++r) g->peer[r] = static_cast<char *>(peer_bases[r]);
    g->local = g->peer[rank];
    g->mc = static_cast<char *>(multicast_base);
    g->owned = false;
    g->connected = true;
    return g;
}

int afk_gather_connect(afk_gather *g, const uint8_t *handles) {
    if (!g || !handles) return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "null argument to afk_gather_connect"));
    cudaStream_t s;
    if (Error e = current_stream(&s, nullptr)) return fail(e);
    for (int r = 0; r < g->world; ++r) {
        if (r == g->rank || g->peer[r]) continue;
        cudaIpcMemHandle_t h;
        memcpy(&h, handles + (size_t)r * 64, 64);
        void *p = nullptr;
        cudaError_t ce = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
        if (ce != cudaSuccess) return fail(cuda_error(ce, "cudaIpcOpenMemHandle (peer result buffer)"));
        g->peer[r] = static_cast<char *>(p);
    }
    g->connected = true;
    return AFK_OK;
}

int afk_gather_band(const afk_gather *g, int rank, int64_t *y0, int64_t *rows) {
    if (!g || rank < 0 || rank >= g->world) return AFK_ERR_UNEXPECTED_INPUT;
    if (y0) *y0 = g->y0[rank];
    if (rows) *rows = g->rows[rank];
    return AFK_OK;
}

// One step on this rank: reduce the band cube [nz, rows(rank), nx] and store the result band into every rank's full
// image (slot = epoch parity), then signal.  Stream-ordered; no host synchronisation.
afk_status afk_gather_reduce(afk_gather *g, const float *band_cube, int64_t nz, int nwin, const int64_t *starts, const int64_t *ends,
                             int epilogue, const float *params, int sum_mode) {
    if (!g || !band_cube || !starts || !ends) return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "null argument to afk_gather_reduce"));
    if (!g->connected) return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "afk_gather_connect has not been called"));
    if (nwin < 1 || nwin > kMaxWindows) return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, fmt("between 1 and %d windows are supported, got %d", kMaxWindows, nwin)));
    const int need = epilogue == EPI_EW ? 3 : (epilogue == EPI_MEAN_RATIO ? 2 : 1);
    if (epilogue < EPI_SUM || epilogue > EPI_EW || nwin != need) return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "epilogue and window count do not match"));
    cudaStream_t s;
    int dev;
    if (Error e = current_stream(&s, &dev)) return fail(e);
    if (dev != g->device) return fail(make_error(AFK_ERR_CUDA, "CUDA: gather belongs to another device"));
    // The two-slot protocol holds only if every reduce is followed by a wait on the same stream before the next reduce:
    // a peer can then not start step n+2 (which overwrites slot n & 1 here) before this rank has finished step n+1,
    // which is ordered after this rank's consumers of image n.
    if (s != g->stream) return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "afk_gather_reduce: called on another stream than the gather was created on"));
    if (g->waited_epoch != g->epoch)
        return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "afk_gather_reduce: afk_gather_wait has not been called since the previous afk_gather_reduce"));
    // every argument check comes before the epoch moves: a rejected call must leave the protocol where it was (a bumped epoch
    // with nothing enqueued would make the next reduce fail and every peer's wait time out)
    ReduceArgs a{};
    a.cube = band_cube;
    a.n2 = g->rows[g->rank] * g->nx;
    a.nwin = nwin;
    for (int w = 0; w < nwin; ++w) {
        if (starts[w] < 0 || !(starts[w] < ends[w]) || ends[w] > nz)
            return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, fmt("Expected 0 <= start < end <= %lld, but got %lld, %lld", (long long)nz, (long long)starts[w], (long long)ends[w])));
        a.start[w] = starts[w];
        a.end[w] = ends[w];
    }
    a.epilogue = epilogue;
    a.mode = sum_mode;
    // params are read for EPI_EW only: exactly 5 floats (a, b, fl, max, sign)
    if (epilogue == EPI_EW && !params) return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "the equivalent-width epilogue needs params[5] = {a, b, fl, max, sign}"));
    for (int i = 0; i < 4; ++i) a.p[i] = epilogue == EPI_EW ? params[i] : 0.0f;
    a.ew_sign = epilogue == EPI_EW ? params[4] : 1.0f;
    // consumers of the previous image on other streams (operators called on its handle from other threads): this step,
    // and therefore every peer's next overwrite of that slot, is ordered after them
    if (afk_image *prev = g->result[g->epoch & 1u]) {
        std::lock_guard<std::mutex> lk(prev->mu);
        for (cudaEvent_t e : prev->readers) {
            cudaStreamWaitEvent(s, e, 0);
            event_put(e);
        }
        prev->readers.clear();
    }
    const unsigned epoch = ++g->epoch;
    const int64_t band_off = g->y0[g->rank] * g->nx;
    int k = 0;
    if (g->mc) { // one multicast store reaches every rank's copy, this rank's included
        a.out = g->mc_slot(epoch) + band_off;
        a.multicast_out = 1;
    } else {
        a.out = g->slot(g->rank, epoch) + band_off;
        for (int r = 0; r < g->world; ++r)
            if (r != g->rank) a.extra_out[k++] = g->slot(r, epoch) + band_off;
    }
    a.n_extra_out = k;
    for (int r = 0; r < g->world; ++r) a.signal[r] = g->flags(r) + g->rank; // my word in rank r's flag array
    a.n_signal = g->world;
    a.done = g->done();
    a.epoch = epoch;
    // min/max exchange: the band's range, found by the kernel's epilogue, travels with the flag to this rank's
    // entry of every rank's range table; afk_gather_wait combines the entries into the gathered image's range memo
    a.range_out = g->range_entry(g->rank, epoch, g->rank);
    k = 0;
    for (int r = 0; r < g->world; ++r)
        if (r != g->rank) a.range_peer[k++] = g->range_entry(r, epoch, g->rank);
    a.n_range_peer = k;
    if (a.n2 <= 0) {
        // an empty band still has to signal: launch a 1-frame plan is impossible, so set the flags from the host side
        // (stream order: the zero range entries land before the flags)
        static thread_local unsigned epoch_src[4]; // outlives the (pageable, hence staged at call time) copies
        epoch_src[epoch & 3u] = epoch;
        for (int r = 0; r < g->world; ++r) {
            cudaError_t ce = cudaMemsetAsync(g->range_entry(r, epoch, g->rank), 0, 4 * sizeof(unsigned), s);
            if (ce != cudaSuccess) return fail(cuda_error(ce, "range entry of an empty band"));
        }
        for (int r = 0; r < g->world; ++r) {
            cudaError_t ce = cudaMemcpyAsync(a.signal[r], &epoch_src[epoch & 3u], sizeof(unsigned), cudaMemcpyHostToDevice, s);
            if (ce != cudaSuccess) return fail(cuda_error(ce, "signal of an empty band"));
        }
        return AFK_OK;
    }
    ReducePlan plan = reduce_plan(a, sm_count(dev));
    ReduceWorkspace ws;
    if (Error e = reduce_workspace(plan, dev, s, &ws)) {
        --g->epoch; // nothing has been enqueued for this epoch
        return fail(e);
    }
    ThreadCtx &c = tls();
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    if (c.profile) {
        cudaEventCreate(&e0);
        cudaEventCreate(&e1);
        cudaEventRecord(e0, s);
    }
    cudaError_t ce = reduce_launch(a, plan, ws, s);
    if (c.profile) {
        cudaEventRecord(e1, s);
        c.prof_events.emplace_back(e0, e1);
        c.prof_launches += 1;
    }
    if (ce != cudaSuccess) return fail(cuda_error(ce, "reduce_kernel launch (fused gather)"));
    return AFK_OK;
}

// Enqueue the wait for the current epoch's bands of all ranks; *result = this rank's full image [ny, nx] of that epoch.
afk_status afk_gather_wait(afk_gather *g, float **result, double timeout_s) {
    if (!g) return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "null argument to afk_gather_wait"));
    cudaStream_t s;
    if (Error e = current_stream(&s, nullptr)) return fail(e);
    if (s != g->stream) return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "afk_gather_wait: called on another stream than the gather was created on"));
    if (g->epoch == 0) return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "afk_gather_wait: no afk_gather_reduce has been enqueued"));
    if (timeout_s <= 0) timeout_s = 5.0;
    // a band that never arrives is overwritten with NaN in this rank's image (and recorded for afk_gather_check), so stale
    // data of an older step cannot flow on silently
    GatherWaitArgs w{};
    w.flags = g->flags(g->rank);
    w.n = g->world;
    w.epoch = g->epoch;
    w.err = g->err();
    w.image = g->slot(g->rank, g->epoch);
    for (int r = 0; r < g->world; ++r) {
        w.band_off[r] = g->y0[r] * g->nx;
        w.band_len[r] = g->rows[r] * g->nx;
    }
    w.ranges = g->range_entry(g->rank, g->epoch, 0);
    w.combined = g->combined(g->epoch);
    cudaError_t ce = gather_wait_launch(w, timeout_s, g->device, s);
    if (ce != cudaSuccess) return fail(cuda_error(ce, "gather_wait_kernel launch"));
    g->waited_epoch = g->epoch;
    if (result) *result = g->slot(g->rank, g->epoch);
    return AFK_OK;
}

// Host-synchronising check: 0 when no wait has timed out, else AFK_ERR_CUDA naming the rank whose band never arrived.
afk_status afk_gather_check(afk_gather *g) {
    if (!g) return AFK_ERR_UNEXPECTED_INPUT;
    cudaStream_t s;
    if (Error e = current_stream(&s, nullptr)) return fail(e);
    unsigned h = 0;
    cudaError_t ce = cudaMemcpyAsync(&h, g->err(), sizeof h, cudaMemcpyDeviceToHost, s);
    if (ce == cudaSuccess) ce = cudaStreamSynchronize(s);
    if (ce != cudaSuccess) return fail(cuda_error(ce, "afk_gather_check"));
    if (h) return fail(make_error(AFK_ERR_CUDA, fmt("CUDA: fused gather timed out waiting for the band of rank %u", h - 1)));
    return AFK_OK;
}

// The handle borrows the slot's memory (never freed through it) and is retained for the caller.  Operators that read it
// from other streams are recorded on it; the next afk_gather_reduce waits for them (see there).
afk_image *afk_gather_result_image(afk_gather *g) {
    if (!g) return nullptr;
    afk_image *&img = g->result[g->epoch & 1u];
    if (!img) {
        int64_t shape[2] = {g->ny, g->nx};
        img = afk_image_wrap_device(g->slot(g->rank, g->epoch), 2, shape);
        if (!img) return nullptr;
        img->track_readers = true;
        // the min/max exchange left the whole image's range beside the flags: image_min_max, logscale and the display chain
        // on this handle make no pass over the 2-D image
        img->range = g->combined(g->epoch);
    }
    img->range_valid = true;
    // ready = everything enqueued on the gather stream so far (the caller has enqueued afk_gather_wait): operators on
    // other streams wait for it before they read
    if (fail(image_mark_ready(img, g->stream)) != AFK_OK) return nullptr;
    afk_image_retain(img);
    return img;
}

// Sharded display chain (aflak_plot/src/imshow/image.rs:20-49): colour-map THIS RANK'S BAND of the gathered image
// of the current epoch -- limits from the whole image's range memo when vmin / vmax (or the chain's logscale range) are NaN
// -- and store the RGBA8 band into every rank's copy of the full [ny, nx, 4] byte image, then signal.  1/world of the
// per-pixel work per GPU; the only traffic is the band's bytes.  Call after afk_gather_wait, on the gather's stream; follow
// with afk_gather_wait_rgba before the next afk_gather_colormap (same two-slot discipline as the f32 result).
afk_status afk_gather_colormap(afk_gather *g, const afk_lut *lut, float vmin, float vmax, int flip_y, const afk_colormap_chain *chain) {
    if (!g || !lut) return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "null argument to afk_gather_colormap"));
    cudaStream_t s;
    int dev;
    if (Error e = current_stream(&s, &dev)) return fail(e);
    if (s != g->stream) return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "afk_gather_colormap: called on another stream than the gather was created on"));
    if (g->epoch == 0 || g->waited_epoch != g->epoch)
        return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "afk_gather_colormap: afk_gather_wait has not been called for the current result"));
    if (g->rgba_waited != g->rgba_epoch)
        return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "afk_gather_colormap: afk_gather_wait_rgba has not been called since the previous afk_gather_colormap"));
    if (g->nx % 4 != 0) return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "afk_gather_colormap: the row length must be a multiple of 4 pixels"));
    const unsigned e2 = g->rgba_epoch + 1; // committed below, once the band has been mapped: a rejected call leaves the protocol where it was
    const int64_t y0 = g->y0[g->rank], rows = g->rows[g->rank];
    const int64_t out_row = flip_y ? g->ny - y0 - rows : y0; // a band of input rows is a band of output rows, mirrored
    const size_t off = (size_t)out_row * (size_t)g->nx * 4;
    afk_status st = AFK_OK;
    if (rows > 0) {
        int64_t shape[2] = {rows, g->nx};
        afk_image *view = afk_image_wrap_device(g->slot(g->rank, g->epoch) + y0 * g->nx, 2, shape);
        if (!view) return AFK_ERR_CUDA;
        view->range = g->combined(g->epoch); // limits are those of the WHOLE image (min/max exchange), not of the band
        view->range_valid = true;
        st = afk_colormap(view, lut, vmin, vmax, flip_y, 4, chain, g->rgba_slot(g->rank, e2) + off);
        afk_image_release(view);
        if (st != AFK_OK) return st;
    }
    g->rgba_epoch = e2;
    unsigned *sig[kMaxPeers];
    for (int r = 0; r < g->world; ++r) sig[r] = g->rgba_flags(r) + g->rank;
    if (rows <= 0) {
        static thread_local unsigned epoch_src[4];
        epoch_src[e2 & 3u] = e2;
        for (int r = 0; r < g->world; ++r) {
            cudaError_t ce = cudaMemcpyAsync(sig[r], &epoch_src[e2 & 3u], sizeof(unsigned), cudaMemcpyHostToDevice, s);
            if (ce != cudaSuccess) return fail(cuda_error(ce, "signal of an empty RGBA band"));
        }
        return AFK_OK;
    }
    BandBroadcastArgs b{};
    b.src = reinterpret_cast<const uint4 *>(g->rgba_slot(g->rank, e2) + off);
    b.n16 = rows * g->nx * 4 / 16;
    int k = 0;
    if (g->mc) {
        b.mc_dst = reinterpret_cast<uint4 *>(g->mc_rgba_slot(e2) + off);
    } else {
        for (int r = 0; r < g->world; ++r)
            if (r != g->rank) b.dst[k++] = reinterpret_cast<uint4 *>(g->rgba_slot(r, e2) + off);
    }
    b.ndst = k;
    b.done = g->rgba_done();
    b.nsig = g->world;
    for (int r = 0; r < g->world; ++r) b.sig[r] = sig[r];
    b.epoch = e2;
    cudaError_t ce = band_broadcast_launch(b, sm_count(dev), s);
    if (ce != cudaSuccess) return fail(cuda_error(ce, "band_broadcast_kernel launch"));
    return AFK_OK;
}

// Enqueue the wait for every rank's RGBA band of the current afk_gather_colormap; *rgba = this rank's full [ny, nx, 4]
// byte image on the device (valid until the second afk_gather_colormap from now).
afk_status afk_gather_wait_rgba(afk_gather *g, uint8_t **rgba, double timeout_s) {
    if (!g) return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "null argument to afk_gather_wait_rgba"));
    cudaStream_t s;
    if (Error e = current_stream(&s, nullptr)) return fail(e);
    if (s != g->stream) return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "afk_gather_wait_rgba: called on another stream than the gather was created on"));
    if (g->rgba_epoch == 0) return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "afk_gather_wait_rgba: no afk_gather_colormap has been enqueued"));
    if (timeout_s <= 0) timeout_s = 5.0;
    GatherWaitArgs w{};
    w.flags = g->rgba_flags(g->rank);
    w.n = g->world;
    w.epoch = g->rgba_epoch;
    w.err = g->err();
    w.image = reinterpret_cast<float *>(g->rgba_slot(g->rank, g->rgba_epoch)); // 4 bytes per pixel: same offsets as the f32 image
    for (int r = 0; r < g->world; ++r) {
        w.band_off[r] = g->y0[r] * g->nx; // a band that never arrives is poisoned where its (unflipped) rows would be
        w.band_len[r] = g->rows[r] * g->nx;
    }
    cudaError_t ce = gather_wait_launch(w, timeout_s, g->device, s);
    if (ce != cudaSuccess) return fail(cuda_error(ce, "gather_wait_kernel launch (RGBA)"));
    g->rgba_waited = g->rgba_epoch;
    if (rgba) *rgba = g->rgba_slot(g->rank, g->rgba_epoch);
    return AFK_OK;
}

void afk_gather_destroy(afk_gather *g) {
    if (!g) return;
    cudaDeviceSynchronize();
    for (afk_image *img : g->result)
        if (img) afk_image_release(img);
    if (g->owned) {
        for (int r = 0; r < g->world; ++r)
            if (r != g->rank && g->peer[r]) cudaIpcCloseMemHandle(g->peer[r]);
        if (g->local) cudaFree(g->local);
    }
    cudaGetLastError();
    delete g;
}

} // extern "C"
