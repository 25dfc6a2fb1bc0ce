// fits.cpp -- FITS -> device ingest: host-side header parsing, the chunked upload + decode pipeline and
// the `open_fits` / `fits_to_image` operators.
//
// Replaces: fitrs 0.5.0 `Fits::open` / `Hdu::read_data` (external crate, Cargo.lock:616; restated from the FITS
// standard: 2880-byte blocks of 80-character cards, big-endian data unit) and aflak_primitives' `WcsArray::from_hdu`
// (unit.rs:283-331), `FitsDataToArray` (fits.rs:42-103), `run_open_fits` / `run_fits_to_image` (lib.rs:826-910).
//
// Data path: file -> pinned double buffer (parallel pread) -> cudaMemcpyAsync -> ingest_kernel (ingest.cu), chunk by
// chunk on two alternating streams, so chunk i is decoded while chunk i+1 crosses PCIe and chunk i+2 is read from the
// file.  4-byte formats land directly in the destination and are decoded in place; 2- and 8-byte formats go through a
// device staging ring.  Only raw file bytes cross PCIe (half the bytes of an f32 upload for BITPIX 16).
#include <errno.h>
#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstring>
#include <thread>

#include "host.h"

namespace afk {

// ---- header --------------------------------------------------------------------------------------------------------
struct FitsCard {
    std::string key;
    char kind = 'U'; // S string, I integer, F real, L logical, U undefined / commentary
    std::string text; // strings: unquoted, right-trimmed; numbers and logicals: the token as written
    int64_t ival = 0;
    double fval = 0.0;
};

struct FitsHdu {
    std::vector<FitsCard> cards;
    int64_t data_offset = 0; // byte offset of the data unit in the file
    int64_t data_bytes = 0;  // unpadded
    int bitpix = 0;
    std::vector<int64_t> naxis; // NAXIS1..NAXISn
    const FitsCard *find(const char *key) const {
        for (const FitsCard &c : cards)
            if (c.key == key) return &c;
        return nullptr;
    }
};

} // namespace afk

struct afk_fits {
    std::atomic<int> rc{1};
    std::string path;
    int fd = -1;
    int64_t file_bytes = 0;
    std::vector<afk::FitsHdu> hdus;
    ~afk_fits() {
        if (fd >= 0) close(fd);
    }
};

namespace afk {

static std::string rtrim(std::string s) {
    while (!s.empty() && (s.back() == ' ' || s.back() == '\0')) s.pop_back();
    return s;
}

// one 80-character card (FITS standard 4.1-4.2, fixed format with free-format tolerance)
static FitsCard parse_card(const char *c) {
    FitsCard card;
    card.key = rtrim(std::string(c, 8));
    if (c[8] != '=' || c[9] != ' ') return card; // commentary (COMMENT, HISTORY, blank) or END
    size_t i = 10;
    while (i < 80 && c[i] == ' ') ++i;
    if (i >= 80) return card;
    if (c[i] == '\'') {
        std::string v;
        ++i;
        while (i < 80) {
            if (c[i] == '\'') {
                if (i + 1 < 80 && c[i + 1] == '\'') {
                    v.push_back('\'');
                    i += 2;
                    continue;
                }
                break;
            }
            v.push_back(c[i++]);
        }
        card.kind = 'S';
        card.text = rtrim(v);
        return card;
    }
    size_t e = i;
    while (e < 80 && c[e] != '/' && c[e] != ' ') ++e;
    std::string tok(c + i, e - i);
    if (tok.empty()) return card;
    card.text = tok;
    if (tok == "T" || tok == "F") {
        card.kind = 'L';
        card.ival = tok == "T";
        return card;
    }
    const bool real = tok.find_first_of(".EeDd") != std::string::npos;
    if (real) {
        std::string t = tok;
        for (char &ch : t)
            if (ch == 'D' || ch == 'd') ch = 'E';
        char *end = nullptr;
        card.fval = strtod(t.c_str(), &end);
        if (end && *end == '\0') card.kind = 'F';
    } else {
        char *end = nullptr;
        errno = 0;
        card.ival = strtoll(tok.c_str(), &end, 10);
        if (end && *end == '\0' && errno == 0) {
            card.kind = 'I';
            card.fval = (double)card.ival;
        }
    }
    return card;
}

static bool hdu_int(const FitsHdu &h, const char *key, int64_t *out) {
    const FitsCard *c = h.find(key);
    if (!c || c->kind != 'I') return false;
    *out = c->ival;
    return true;
}
// read_float (unit.rs:260-268): a real, or an integer widened to f64
static bool hdu_float(const FitsHdu &h, const char *key, double *out) {
    const FitsCard *c = h.find(key);
    if (!c || (c->kind != 'F' && c->kind != 'I')) return false;
    *out = c->fval;
    return true;
}
// read_string (unit.rs:252-258)
static bool hdu_string(const FitsHdu &h, const char *key, std::string *out) {
    const FitsCard *c = h.find(key);
    if (!c || c->kind != 'S') return false;
    *out = c->text;
    return true;
}

static Error parse_file(afk_fits *f) {
    int64_t off = 0;
    std::vector<char> block(2880);
    while (off + 2880 <= f->file_bytes) {
        FitsHdu h;
        bool end = false;
        while (!end) {
            if (off + 2880 > f->file_bytes) return make_error(AFK_ERR_FITS, "Unexpected end of file inside a header");
            if (pread(f->fd, block.data(), 2880, off) != 2880) return make_error(AFK_ERR_FITS, "Short read inside a header");
            off += 2880;
            if (h.cards.empty()) { // first block of this HDU: it must open with SIMPLE (primary) or XTENSION
                const bool primary = f->hdus.empty();
                if (memcmp(block.data(), primary ? "SIMPLE  " : "XTENSION", 8) != 0) {
                    if (primary) return make_error(AFK_ERR_FITS, "Not a FITS file (the first card is not SIMPLE)");
                    return ok(); // trailing bytes after the last HDU
                }
            }
            for (int i = 0; i < 36; ++i) {
                const char *c = block.data() + 80 * i;
                if (memcmp(c, "END     ", 8) == 0) {
                    end = true;
                    break;
                }
                h.cards.push_back(parse_card(c));
            }
        }
        int64_t bitpix = 0, naxis = 0;
        if (!hdu_int(h, "BITPIX", &bitpix) || !hdu_int(h, "NAXIS", &naxis) || naxis < 0 || naxis > 999)
            return make_error(AFK_ERR_FITS, "Header without a valid BITPIX / NAXIS");
        h.bitpix = (int)bitpix;
        int64_t count = naxis > 0 ? 1 : 0;
        for (int64_t a = 1; a <= naxis; ++a) {
            int64_t n = 0;
            if (!hdu_int(h, fmt("NAXIS%lld", (long long)a).c_str(), &n) || n < 0) return make_error(AFK_ERR_FITS, "Header without a valid NAXISn");
            h.naxis.push_back(n);
            if (__builtin_mul_overflow(count, n, &count)) return make_error(AFK_ERR_FITS, "Header whose NAXISn product overflows");
        }
        if (bitpix != 8 && bitpix != 16 && bitpix != 32 && bitpix != 64 && bitpix != -32 && bitpix != -64)
            return make_error(AFK_ERR_FITS, "Header with an invalid BITPIX");
        int64_t pcount = 0, gcount = 1;
        hdu_int(h, "PCOUNT", &pcount);
        hdu_int(h, "GCOUNT", &gcount);
        if (pcount < 0 || gcount < 1) return make_error(AFK_ERR_FITS, "Header with a negative PCOUNT or a GCOUNT below 1");
        h.data_offset = off;
        // data unit size = |BITPIX|/8 * GCOUNT * (PCOUNT + NAXIS1*...*NAXISn), checked: the header is untrusted input
        int64_t elems = 0, bytes = 0, padded = 0;
        if (__builtin_add_overflow(pcount, count, &elems) || __builtin_mul_overflow(elems, gcount, &elems) ||
            __builtin_mul_overflow(elems, std::llabs(bitpix) / 8, &bytes) || __builtin_add_overflow(bytes, (int64_t)2879, &padded))
            return make_error(AFK_ERR_FITS, "Header whose data unit size overflows");
        padded = padded / 2880 * 2880;
        h.data_bytes = bytes;
        // a data unit that does not fit in the file ends the HDU list (`off` always advances, so the loop terminates)
        const bool truncated = bytes > f->file_bytes - off;
        off += padded;
        f->hdus.push_back(std::move(h));
        if (truncated) break; // the header stays readable; reading the data unit fails in hdu_job (bounds check there)
    }
    if (f->hdus.empty()) return make_error(AFK_ERR_FITS, "Not a FITS file (shorter than one block)");
    return ok();
}

// ---- upload + decode pipeline ----------------------------------------------------------------------------------------
struct IngestStats {
    double read_ms = 0, wall_ms = 0, device_ms = 0;
    uint64_t bytes = 0;
};
static thread_local IngestStats t_stats;
static thread_local cudaEvent_t t_ingest_ev0 = nullptr, t_ingest_ev1 = nullptr; // device span of the last ingest

struct IngestJob {
    int fmt = ING_F32;
    int has_blank = 0;
    int64_t blank = 0;
    float bscale = 1.0f, bzero = 0.0f;
    // geometry: nz frames of ny x nx elements in the source; rows [y0, y0+rows) of every frame are taken
    int64_t nz = 1, ny = 1, nx = 1, y0 = 0, rows = 1;
    // source: memory (raw != nullptr) or file (fd >= 0, data unit at file_off)
    const uint8_t *raw = nullptr;
    int fd = -1;
    int64_t file_off = 0;
};

constexpr size_t kIngestChunkBytes = (size_t)64 << 20; // raw bytes per pipeline stage
constexpr int kReaderThreads = 8;

struct PinnedRing { // per-thread pinned double buffer for file sources
    void *buf[2] = {nullptr, nullptr};
    size_t bytes = 0;
    cudaEvent_t free_ev[2] = {nullptr, nullptr}; // recorded after the H2D copy that last read buf[i]
    ~PinnedRing() {
        for (int i = 0; i < 2; ++i) {
            if (buf[i]) cudaFreeHost(buf[i]);
            if (free_ev[i]) cudaEventDestroy(free_ev[i]);
        }
    }
};
static thread_local PinnedRing t_ring;

static Error ring_reserve(size_t bytes) {
    if (t_ring.bytes >= bytes) return ok();
    for (int i = 0; i < 2; ++i) {
        if (t_ring.buf[i]) cudaFreeHost(t_ring.buf[i]);
        t_ring.buf[i] = nullptr;
        AFK_CUDA(cudaHostAlloc(&t_ring.buf[i], bytes, cudaHostAllocDefault));
        if (!t_ring.free_ev[i]) AFK_CUDA(cudaEventCreateWithFlags(&t_ring.free_ev[i], cudaEventDisableTiming));
    }
    t_ring.bytes = bytes;
    return ok();
}

struct IngestStreams {
    cudaStream_t s[2] = {nullptr, nullptr};
    int device = -1;
};
static thread_local IngestStreams t_ing_streams[16];

static Error ingest_streams(int dev, cudaStream_t out[2]) {
    IngestStreams &st = t_ing_streams[dev];
    for (int i = 0; i < 2; ++i) {
        if (!st.s[i]) AFK_CUDA(cudaStreamCreateWithFlags(&st.s[i], cudaStreamNonBlocking));
        out[i] = st.s[i];
    }
    return ok();
}

// Read one pipeline chunk into pinned memory with several threads (page cache -> pinned memory is a memcpy; one thread
// does ~5-10 GB/s).  The chunk is `frames` segments of seg_bytes each, segment k at file offset off0 + k * seg_stride
// (a contiguous chunk is one segment): the threads share the total byte range evenly, whatever the segment size.
static bool read_chunk(int fd, uint8_t *dst, int64_t frames, size_t seg_bytes, int64_t off0, int64_t seg_stride) {
    const size_t total = (size_t)frames * seg_bytes;
    auto rd = [=](size_t b, size_t e) { // bytes [b, e) of the chunk
        while (b < e) {
            const size_t k = b / seg_bytes, in = b - k * seg_bytes;
            size_t n = std::min(e - b, seg_bytes - in);
            int64_t o = off0 + (int64_t)k * seg_stride + (int64_t)in;
            uint8_t *d = dst + b;
            b += n;
            while (n > 0) {
                const ssize_t r = pread(fd, d, n, o);
                if (r <= 0) return false;
                d += r;
                o += r;
                n -= (size_t)r;
            }
        }
        return true;
    };
    if (total < ((size_t)4 << 20)) return rd(0, total);
    bool okv[kReaderThreads];
    std::thread th[kReaderThreads];
    const size_t per = (total / kReaderThreads + 4095) & ~(size_t)4095;
    for (int t = 0; t < kReaderThreads; ++t) {
        const size_t lo = std::min(total, per * t), hi = std::min(total, per * (t + 1));
        okv[t] = true;
        th[t] = std::thread([&, t, lo, hi] { okv[t] = hi > lo ? rd(lo, hi) : true; });
    }
    bool good = true;
    for (int t = 0; t < kReaderThreads; ++t) {
        th[t].join();
        good = good && okv[t];
    }
    return good;
}

// Runs the pipeline into `dst` ([nz, rows, nx] f32, already allocated on stream `s`).
static Error ingest_run(const IngestJob &j, float *dst, cudaStream_t s, int dev) {
    const auto t_begin = std::chrono::steady_clock::now();
    const int es = ingest_element_bytes(j.fmt);
    if (es == 0) return make_error(AFK_ERR_FITS, "Unsupported data array format");
    const int64_t band_elems = j.rows * j.nx;           // elements taken from one frame
    const int64_t total = j.nz * band_elems;
    t_stats = IngestStats();
    t_stats.bytes = (uint64_t)total * (uint64_t)es;
    if (total <= 0) return ok();
    const bool full_frames = (j.y0 == 0 && j.rows == j.ny);
    // a chunk is a whole number of frames (band mode) or an arbitrary element range (contiguous mode)
    int64_t chunk_elems;
    if (full_frames) chunk_elems = (int64_t)(kIngestChunkBytes / (size_t)es) & ~(int64_t)3;
    else {
        int64_t frames = (int64_t)(kIngestChunkBytes / (size_t)es) / band_elems;
        if (frames < 1) frames = 1;
        chunk_elems = frames * band_elems;
    }
    const size_t chunk_bytes = (size_t)chunk_elems * (size_t)es;
    cudaStream_t st[2];
    if (Error e = ingest_streams(dev, st)) return e;
    const bool in_place = (es == 4);
    void *stage[2] = {nullptr, nullptr};
    Error err = ok();
    if (j.fd >= 0) err = ring_reserve(chunk_bytes);
    if (err) return err;
    // fork: the two pipeline streams start after everything already queued on s (the allocation of dst and stage)
    cudaEvent_t fork = event_get(), ev_t0 = nullptr, ev_t1 = nullptr;
    cudaEventCreate(&ev_t0);
    cudaEventCreate(&ev_t1);
    {
        cudaError_t ce = cudaSuccess;
        if (!in_place)
            for (int i = 0; i < 2 && ce == cudaSuccess; ++i) ce = cudaMallocAsync(&stage[i], chunk_bytes, s);
        if (ce == cudaSuccess) ce = cudaEventRecord(fork, s);
        if (ce == cudaSuccess) ce = cudaEventRecord(ev_t0, s);
        for (int i = 0; i < 2 && ce == cudaSuccess; ++i) ce = cudaStreamWaitEvent(st[i], fork, 0);
        if (ce != cudaSuccess) err = cuda_error(ce, "ingest pipeline set-up"); // falls through to the clean-up below
    }
    int64_t done = 0;
    for (int c = 0; done < total && !err; ++c) {
        const int b = c & 1;
        const int64_t n = std::min(chunk_elems, total - done);
        const size_t nbytes = (size_t)n * (size_t)es;
        uint8_t *dev_raw = in_place ? reinterpret_cast<uint8_t *>(dst + done) : static_cast<uint8_t *>(stage[b]);
        const uint8_t *host_src = nullptr;
        bool strided = false; // 2-D copy straight from user memory
        int64_t k0 = 0, frames = 0;
        if (!full_frames) {
            k0 = done / band_elems;
            frames = n / band_elems;
        }
        if (j.fd >= 0) {
            // the pinned buffer is free once the copy issued two chunks ago has drained
            if (c >= 2) {
                const cudaError_t we = cudaEventSynchronize(t_ring.free_ev[b]);
                if (we != cudaSuccess) {
                    err = cuda_error(we, "cudaEventSynchronize(pinned ring)");
                    break;
                }
            }
            const auto r0 = std::chrono::steady_clock::now();
            uint8_t *pin = static_cast<uint8_t *>(t_ring.buf[b]);
            const bool good = full_frames ? read_chunk(j.fd, pin, 1, nbytes, j.file_off + done * es, 0)
                                          : read_chunk(j.fd, pin, frames, (size_t)band_elems * es,
                                                       j.file_off + ((k0 * j.ny + j.y0) * j.nx) * es, j.ny * j.nx * es);
            t_stats.read_ms += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - r0).count();
            if (!good) {
                err = make_error(AFK_ERR_FITS, "Unexpected end of file inside the data unit");
                break;
            }
            host_src = pin;
        } else if (full_frames) {
            host_src = j.raw + (size_t)done * es;
        } else {
            host_src = j.raw + (size_t)((k0 * j.ny + j.y0) * j.nx) * es;
            strided = true;
        }
        cudaError_t ce;
        if (strided)
            ce = cudaMemcpy2DAsync(dev_raw, (size_t)band_elems * es, host_src, (size_t)j.ny * j.nx * es, (size_t)band_elems * es,
                                   (size_t)frames, cudaMemcpyHostToDevice, st[b]);
        else ce = cudaMemcpyAsync(dev_raw, host_src, nbytes, cudaMemcpyHostToDevice, st[b]);
        if (ce == cudaSuccess && j.fd >= 0) ce = cudaEventRecord(t_ring.free_ev[b], st[b]);
        if (ce == cudaSuccess) {
            IngestArgs a{dev_raw, j.fmt, n, j.has_blank, j.blank, j.bscale, j.bzero, dst + done};
            ce = ingest_launch(a, sm_count(dev), st[b]);
        }
        if (ce != cudaSuccess) {
            err = cuda_error(ce, "ingest pipeline");
            break;
        }
        done += n;
    }
    // join: s continues after both pipeline streams
    for (int i = 0; i < 2; ++i) {
        cudaEventRecord(fork, st[i]);
        cudaStreamWaitEvent(s, fork, 0);
    }
    cudaEventRecord(ev_t1, s);
    for (int i = 0; i < 2; ++i)
        if (stage[i]) cudaFreeAsync(stage[i], s);
    event_put(fork);
    if (!err && j.fd >= 0) {
        // the pinned ring is reused by the next call on this thread: wait for the last copies (the decode may still run)
        cudaEventSynchronize(t_ring.free_ev[0]);
        cudaEventSynchronize(t_ring.free_ev[1]);
    }
    t_stats.wall_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_begin).count();
    t_stats.device_ms = -1.0; // filled by afk_ingest_last_stats once the stream has drained
    if (t_ingest_ev0) cudaEventDestroy(t_ingest_ev0);
    if (t_ingest_ev1) cudaEventDestroy(t_ingest_ev1);
    t_ingest_ev0 = ev_t0;
    t_ingest_ev1 = ev_t1;
    return err;
}

static Error ingest_image(const IngestJob &j, int ndim, const int64_t *shape, afk_image **out) {
    afk_image *img = nullptr;
    if (Error e = image_new(ndim, shape, &img)) return e;
    int dev = img->device;
    Error err = ingest_run(j, img->ptr, img->stream, dev);
    if (!err) err = image_mark_ready(img, img->stream);
    if (err) {
        afk_image_release(img);
        return err;
    }
    *out = img;
    return ok();
}

// ---- WcsArray::from_hdu (unit.rs:283-331) ---------------------------------------------------------------------------------
static Error hdu_to_image(const afk_fits *f, const FitsHdu &h, int world, int rank, afk_image **out) {
    IngestJob j;
    switch (h.bitpix) {
    case -32: j.fmt = ING_F32; break;
    case -64: j.fmt = ING_F64; break;
    case 16: j.fmt = ING_I16; break;
    case 32: j.fmt = ING_I32; break;
    case 8: return make_error(AFK_ERR_FITS, "Unsupported data array format: 'Characters'."); // FitsArrayReadError::UnsupportedData
    default: return make_error(AFK_ERR_FITS, fmt("Unsupported BITPIX %d", h.bitpix));
    }
    const int ndim = (int)h.naxis.size();
    if (ndim < 1) return make_error(AFK_ERR_FITS, "ShapeError/IncompatibleShape: the HDU has no data (NAXIS = 0)");
    if (ndim > AFK_MAX_NDIM) return make_error(AFK_ERR_FITS, fmt("FITS arrays of more than %d dimensions are not supported (NAXIS = %d)", AFK_MAX_NDIM, ndim));
    int64_t shape[AFK_MAX_NDIM]; // [NAXISn..NAXIS1] (fits.rs:46)
    int64_t count = 1;
    for (int i = 0; i < ndim; ++i) {
        shape[i] = h.naxis[ndim - 1 - i];
        count *= shape[i];
    }
    if (count * ingest_element_bytes(j.fmt) > h.data_bytes || h.data_offset + h.data_bytes > f->file_bytes)
        return make_error(AFK_ERR_FITS, "Unexpected end of file inside the data unit");
    int64_t blank = 0;
    j.has_blank = (j.fmt == ING_I16 || j.fmt == ING_I32) && hdu_int(h, "BLANK", &blank);
    j.blank = blank;
    double bz = 0.0, bs = 1.0; // unit.rs:306-318: absent -> 0.0 / 1.0, present -> `as f32`
    if (hdu_float(h, "BZERO", &bz)) j.bzero = (float)bz;
    if (hdu_float(h, "BSCALE", &bs)) j.bscale = (float)bs;
    j.fd = f->fd;
    j.file_off = h.data_offset;
    // frames x rows x columns view of the array for the band split: axis 0 of the array is the frame axis
    if (world > 1) {
        if (ndim != 3) return unexpected(fmt("A sharded upload needs a 3-dimensional HDU, got %d dimensions", ndim));
        j.nz = shape[0];
        j.ny = shape[1];
        j.nx = shape[2];
        if (afk_shard_band(j.ny, world, rank, &j.y0, &j.rows) != AFK_OK) return unexpected("bad world / rank for a sharded upload");
        shape[1] = j.rows;
    } else {
        j.nz = 1;
        j.ny = 1;
        j.nx = count;
        j.y0 = 0;
        j.rows = 1;
    }
    afk_image *img = nullptr;
    if (Error e = ingest_image(j, ndim, shape, &img)) return e;
    ImageMeta &m = img->meta;
    m.has_meta = true;
    m.has_unit = hdu_string(h, "BUNIT", &m.unit);
    m.homogeneous = true;
    m.wcs.assign(12, 0.0);
    for (int a = 0; a < 4; ++a) {
        AxisMeta &ax = m.axes[a];
        ax.has_name = hdu_string(h, fmt("CTYPE%d", a + 1).c_str(), &ax.name);
        if (!hdu_string(h, fmt("CUNIT%d", a + 1).c_str(), &ax.unit)) ax.unit.clear();
        double crval = 0.0, crpix = 1.0, cdelt = 1.0; // absent keys: the axis maps a pixel to itself
        hdu_float(h, fmt("CRVAL%d", a + 1).c_str(), &crval);
        hdu_float(h, fmt("CRPIX%d", a + 1).c_str(), &crpix);
        hdu_float(h, fmt("CDELT%d", a + 1).c_str(), &cdelt);
        m.wcs[3 * a + 0] = crval;
        m.wcs[3 * a + 1] = crpix;
        m.wcs[3 * a + 2] = cdelt;
    }
    if (world > 1) m.wcs[3 * 1 + 1] -= (double)j.y0; // the band starts at row y0: shift CRPIX2 so world coordinates are unchanged
    *out = img;
    return ok();
}

static Error find_hdu(const afk_fits *f, int64_t hdu_idx, const char *extension, const FitsHdu **out) {
    if (Error e = try_into_unsigned("hdu_idx", hdu_idx)) return e;
    const std::string ext = extension ? extension : "";
    // fits.get_by_name(extension).or_else(|| fits.get(hdu_idx)) (lib.rs:889-891)
    for (const FitsHdu &h : f->hdus) {
        std::string name;
        if (hdu_string(h, "EXTNAME", &name) && name == ext) {
            *out = &h;
            return ok();
        }
    }
    if ((size_t)hdu_idx < f->hdus.size()) {
        *out = &f->hdus[(size_t)hdu_idx];
        return ok();
    }
    const std::string hdu_name = hdu_idx == 0 ? "Primary HDU" : fmt("HDU #%lld", (long long)hdu_idx);
    if (ext.empty()) return unexpected(fmt("Could not find %s in FITS file.", hdu_name.c_str()));
    return unexpected(fmt("Could not find HDU '%s', nor %s in FITS file.", ext.c_str(), hdu_name.c_str()));
}

static std::string rust_path_debug(const std::string &p) { // {:?} of a Path: quoted, with \ and " escaped
    std::string s = "\"";
    for (char c : p) {
        if (c == '"' || c == '\\') s.push_back('\\');
        s.push_back(c);
    }
    return s + "\"";
}

static Error fits_open(const char *path, afk_fits **out) {
    const std::string p = path ? path : "";
    const int fd = open(p.c_str(), O_RDONLY | O_CLOEXEC);
    if (fd < 0) {
        const int en = errno;
        // IOErr::IoError(err, "Could not open file {:?}") displayed as "{}. This was caused by '{}'." (lib.rs:122)
        return make_error(AFK_ERR_IO, fmt("Could not open file %s. This was caused by '%s (os error %d)'", rust_path_debug(p).c_str(), strerror(en), en));
    }
    afk_fits *f = new afk_fits();
    f->path = p;
    f->fd = fd;
    struct stat sb;
    if (fstat(fd, &sb) == 0) f->file_bytes = (int64_t)sb.st_size;
    if (Error e = parse_file(f)) {
        delete f;
        return make_error(AFK_ERR_IO, fmt("Could not open file %s. This was caused by '%s'", rust_path_debug(p).c_str(), e.message.c_str()));
    }
    *out = f;
    return ok();
}

// ---- operators -------------------------------------------------------------------------------------------------------
afk_status op_open_fits(const afk_value *const *in, size_t n_in, afk_result *out, size_t n_out) { // lib.rs:145-153, 826-835
    if (n_out != 1) return fail(make_error(AFK_ERR_PANIC, "wrong number of output slots"));
    if (!args_match(in, n_in, {AFK_T_PATHS, AFK_T_INTEGER}, out, n_out)) return AFK_ERR_PANIC;
    const std::vector<std::string> &paths = in[0]->paths;
    const uint64_t n = (uint64_t)in[1]->i; // `n as usize`
    Error e = ok();
    if (!((uint64_t)paths.size() > n)) e = unexpected(fmt("Expected pathlist_len > n, but got %zu > %llu", paths.size(), (unsigned long long)n));
    afk_fits *f = nullptr;
    if (!e) e = fits_open(paths[(size_t)n].c_str(), &f);
    if (e) {
        set_err(&out[0], e);
        return e.status;
    }
    afk_value *v = afk_value_fits(f);
    afk_fits_release(f);
    set_ok(&out[0], v);
    return AFK_OK;
}

afk_status op_fits_to_image(const afk_value *const *in, size_t n_in, afk_result *out, size_t n_out) { // lib.rs:163-170, 883-910
    if (n_out != 1) return fail(make_error(AFK_ERR_PANIC, "wrong number of output slots"));
    if (!args_match(in, n_in, {AFK_T_FITS, AFK_T_INTEGER, AFK_T_STR}, out, n_out)) return AFK_ERR_PANIC;
    afk_image *img = nullptr;
    const afk_status st = afk_fits_to_image(in[0]->fits, in[1]->i, in[2]->s.c_str(), &img);
    if (st != AFK_OK) {
        set_err(&out[0], make_error(st, afk_last_error()));
        return st;
    }
    set_image(&out[0], img);
    return AFK_OK;
}

} // namespace afk

using namespace afk;

extern "C" {

afk_fits *afk_fits_open(const char *path) {
    afk_fits *f = nullptr;
    if (fail(fits_open(path, &f)) != AFK_OK) return nullptr;
    return f;
}
void afk_fits_retain(afk_fits *f) {
    if (f) f->rc.fetch_add(1, std::memory_order_relaxed);
}
void afk_fits_release(afk_fits *f) {
    if (f && f->rc.fetch_sub(1, std::memory_order_acq_rel) == 1) delete f;
}
size_t afk_fits_hdu_count(const afk_fits *f) { return f ? f->hdus.size() : 0; }

const char *afk_fits_header(const afk_fits *f, size_t hdu, const char *key) {
    if (!f || !key || hdu >= f->hdus.size()) return nullptr;
    const FitsCard *c = f->hdus[hdu].find(key);
    return (c && c->kind != 'U') ? c->text.c_str() : nullptr;
}

afk_status afk_fits_to_image_band(const afk_fits *f, int64_t hdu_idx, const char *extension, int world, int rank, afk_image **out) {
    if (!f || !out) return fail(make_error(AFK_ERR_PANIC, "Unexpected argument!"));
    const FitsHdu *h = nullptr;
    if (Error e = find_hdu(f, hdu_idx, extension, &h)) return fail(e);
    return fail(hdu_to_image(f, *h, world, rank, out));
}

afk_status afk_fits_to_image(const afk_fits *f, int64_t hdu_idx, const char *extension, afk_image **out) {
    return afk_fits_to_image_band(f, hdu_idx, extension, 1, 0, out);
}

afk_image *afk_image_ingest_band(const void *raw_be, afk_fits_format fmt_, int64_t nz, int64_t ny, int64_t nx, int64_t y0, int64_t rows,
                                 int has_blank, int64_t blank, float bscale, float bzero) {
    if (!raw_be || ingest_element_bytes((int)fmt_) == 0 || nz < 0 || ny < 0 || nx < 0 || y0 < 0 || rows < 0 || y0 + rows > ny) {
        fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "bad arguments to afk_image_ingest_band"));
        return nullptr;
    }
    IngestJob j;
    j.fmt = (int)fmt_;
    j.has_blank = has_blank && (fmt_ == AFK_FITS_I16 || fmt_ == AFK_FITS_I32 || fmt_ == AFK_FITS_U32);
    j.blank = blank;
    j.bscale = bscale;
    j.bzero = bzero;
    j.nz = nz; j.ny = ny; j.nx = nx; j.y0 = y0; j.rows = rows;
    j.raw = static_cast<const uint8_t *>(raw_be);
    const int64_t shape[3] = {nz, rows, nx};
    afk_image *img = nullptr;
    if (fail(ingest_image(j, 3, shape, &img)) != AFK_OK) return nullptr;
    return img;
}

afk_image *afk_image_ingest(const void *raw_be, afk_fits_format fmt_, int ndim, const int64_t *shape, int has_blank, int64_t blank,
                            float bscale, float bzero) {
    if (!raw_be || ingest_element_bytes((int)fmt_) == 0 || ndim < 0 || ndim > AFK_MAX_NDIM || (ndim > 0 && !shape)) {
        fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "bad arguments to afk_image_ingest"));
        return nullptr;
    }
    IngestJob j;
    j.fmt = (int)fmt_;
    j.has_blank = has_blank && (fmt_ == AFK_FITS_I16 || fmt_ == AFK_FITS_I32 || fmt_ == AFK_FITS_U32);
    j.blank = blank;
    j.bscale = bscale;
    j.bzero = bzero;
    int64_t count = 1;
    for (int i = 0; i < ndim; ++i) count *= shape[i];
    j.nx = count;
    j.raw = static_cast<const uint8_t *>(raw_be);
    afk_image *img = nullptr;
    if (fail(ingest_image(j, ndim, shape, &img)) != AFK_OK) return nullptr;
    return img;
}

afk_status afk_k_ingest(const void *device_raw_be, afk_fits_format fmt_, int64_t n, int has_blank, int64_t blank, float bscale, float bzero,
                        float *device_out) {
    cudaStream_t s;
    int dev;
    if (Error e = current_stream(&s, &dev)) return fail(e);
    IngestArgs a{device_raw_be, (int)fmt_, n, has_blank, blank, bscale, bzero, device_out};
    const cudaError_t ce = ingest_launch(a, sm_count(dev), s);
    if (ce != cudaSuccess) return fail(cuda_error(ce, "ingest_kernel launch"));
    return AFK_OK;
}

int afk_ingest_last_stats(double ms[3], uint64_t *bytes) {
    if (!ms) return AFK_ERR_UNEXPECTED_INPUT;
    if (t_stats.device_ms < 0 && t_ingest_ev0 && t_ingest_ev1) {
        float d = 0.0f;
        if (cudaEventSynchronize(t_ingest_ev1) == cudaSuccess && cudaEventElapsedTime(&d, t_ingest_ev0, t_ingest_ev1) == cudaSuccess)
            t_stats.device_ms = d;
        else cudaGetLastError();
    }
    ms[0] = t_stats.read_ms;
    ms[1] = t_stats.wall_ms;
    ms[2] = t_stats.device_ms;
    if (bytes) *bytes = t_stats.bytes;
    return AFK_OK;
}

afk_value *afk_value_fits(afk_fits *f) {
    if (!f) return nullptr;
    afk_value *v = new afk_value();
    v->type = AFK_T_FITS;
    afk_fits_retain(f);
    v->fits = f;
    return v;
}
afk_fits *afk_value_get_fits(const afk_value *v) { return (v && v->type == AFK_T_FITS) ? v->fits : nullptr; }

afk_value *afk_value_paths(const char *const *paths, size_t n) {
    afk_value *v = new afk_value();
    v->type = AFK_T_PATHS;
    for (size_t i = 0; i < n; ++i) v->paths.emplace_back(paths && paths[i] ? paths[i] : "");
    return v;
}
size_t afk_value_paths_count(const afk_value *v) { return (v && v->type == AFK_T_PATHS) ? v->paths.size() : 0; }
const char *afk_value_paths_at(const afk_value *v, size_t i) {
    return (v && v->type == AFK_T_PATHS && i < v->paths.size()) ? v->paths[i].c_str() : nullptr;
}

} // extern "C"
