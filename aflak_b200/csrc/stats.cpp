// stats.cpp -- host side of the display statistics: histogram, normalised median / MAD, zscale.
//
// Replaces aflak_plot/src/imshow/hist.rs:10-66, lims.rs:37-106 and zscale.rs:45-242 (`Image::new` runs the first two on
// every new image, `zscale` on demand; all on the UI thread).  The image passes run on the device (stats.cu); zscale's
// line fit works on <= ~1000 samples and stays on the host, restated step by step in f32.
#include <algorithm>
#include <cmath>
#include <cstddef>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <vector>

#include "host.h"

namespace afk {

static Error stats_ctx(const afk_image *img, cudaStream_t *s, int *dev) {
    if (Error e = current_stream(s, dev)) return e;
    if (img->device != *dev)
        return make_error(AFK_ERR_CUDA, fmt("CUDA: image lives on device %d but the calling thread uses device %d", img->device, *dev));
    return image_wait(img, *s);
}

static Error histogram_run(const afk_image *img, int channels, int nbins, float vmin, float vmax, uint64_t *counts, float *starts, float *ends) {
    cudaStream_t s;
    int dev;
    if (Error e = stats_ctx(img, &s, &dev)) return e;
    const int64_t n_per = img->len() / channels;
    const size_t entries = (size_t)nbins * channels, bytes = entries * sizeof(unsigned long long);
    if (entries > kHistSelfEntries) return make_error(AFK_ERR_UNEXPECTED_INPUT, "histogram: too many bins");
    // The counts reach the host through this thread's mapped read-back page: the kernel's last CTA stores them there and
    // returns the device block to zero, so the call is one launch and one synchronisation (no allocation, memset or copy).
    char *page = nullptr;
    if (Error e = readback_page(&page)) return e;
    unsigned long long *host_out = reinterpret_cast<unsigned long long *>(page + 1024);
    memset(host_out, 0, bytes); // an empty image launches nothing
    StreamScratch &sc = tls().scratch[std::make_pair(dev, s)];
    cudaError_t ce = cudaSuccess;
    if (!sc.hist_self) {
        const size_t block = (kHistSelfEntries + 2) * sizeof(unsigned long long); // counts, then the ticket word
        ce = cudaMallocAsync(reinterpret_cast<void **>(&sc.hist_self), block, s);
        if (ce == cudaSuccess) ce = cudaMemsetAsync(sc.hist_self, 0, block, s);
        else sc.hist_self = nullptr;
    }
    if (ce == cudaSuccess)
        ce = histogram_launch(img->ptr, n_per, channels, vmin, vmax, nbins, sc.hist_self, sm_count(dev), s,
                              reinterpret_cast<unsigned *>(sc.hist_self + kHistSelfEntries), host_out);
    if (ce == cudaSuccess) ce = cudaStreamSynchronize(s);
    if (ce != cudaSuccess) {
        if (sc.hist_self) cudaFreeAsync(sc.hist_self, s); // the block may not have been returned to zero
        sc.hist_self = nullptr;
        return cuda_error(ce, "histogram_kernel");
    }
    memcpy(counts, host_out, bytes);
    image_note_read(img, s);
    // Bin { start: min + i as f32 * size, end: min + (i + 1) as f32 * size } with size = (max - min) / NB (hist.rs:16-23)
    const float size = (vmax - vmin) / (float)nbins;
    for (int i = 0; i < nbins; ++i) {
        if (starts) starts[i] = vmin + (float)i * size;
        if (ends) ends[i] = vmin + (float)(i + 1) * size;
    }
    return ok();
}

static Error vmed_vmad_run(const afk_image *img, float *vmed, float *vmad) {
    if (img->len() == 0) return make_error(AFK_ERR_UNEXPECTED_INPUT, "Could not get vmed from normalized image"); // get_vmax/get_vmin fail
    cudaStream_t s;
    int dev;
    if (Error e = stats_ctx(img, &s, &dev)) return e;
    const int64_t n = img->len();
    static const int one_pass_env = getenv("AFK_STATS_ONE_PASS") ? atoi(getenv("AFK_STATS_ONE_PASS")) : 1; // 0: radix select only (A/B)
    const bool one_pass = one_pass_env != 0 && n >= kStatBracketMinPixels;
    // get_vmin / get_vmax (lims.rs:7-35).  The radix select needs them on the host before its first launch; the one-pass select
    // leaves them on the device (its kernels decode the words and decide the raw-key shortcut themselves: stat_resolve), so the
    // whole chain has ONE host synchronisation, at its end.
    float vmin = 0.0f, vmax = 0.0f;
    bool any = false;
    if (!one_pass)
        if (Error e = op_min_max(img, &vmin, &vmax, &any)) return e;
    // device block: the states of both selections (one-pass bracket select and radix select), their histograms, the small
    // buffers of the bracket finish, then the candidate buffer the selections share (they run one after the other)
    struct Head {
        // read back in ONE copy at the end of the one-pass chain (sel, br, mm are contiguous): a pageable read-back is a host
        // synchronisation of its own, three of them cost ~15 us of a ~200 us chain
        StatState sel[2];     // results of the one-pass select (vmed, vmad)
        StatBracket br[2];
        unsigned mm[4];       // minmax_kernel's words for the one-pass select (when the image carries no range memo)
        StatState radix[2];   // radix select: the path for small images and the fallback
        unsigned hist_radix[256];
        unsigned hist_br[2][kStatBracketBins];
        unsigned binbuf[2][kStatBracketBinBuf];
    };
    const size_t head = (offsetof(Head, binbuf) + 15) / 16 * 16; // initialised from the host up to here
    const unsigned cand_cap = (unsigned)(n < (int64_t(1) << 22) ? n : (int64_t(1) << 22)); // 4 Mi keys: 25 % of a 4096^2 image
    char *d = nullptr;
    AFK_CUDA(cudaMallocAsync(reinterpret_cast<void **>(&d), sizeof(Head) + (size_t)cand_cap * sizeof(unsigned), s));
    Head *hd = reinterpret_cast<Head *>(d);
    unsigned *cand = reinterpret_cast<unsigned *>(d + sizeof(Head));
    static const std::vector<char> init_block = [head] {
        std::vector<char> b(head, 0);
        Head *h0 = reinterpret_cast<Head *>(b.data()); // only the fields below `binbuf` are touched
        for (int i = 0; i < 2; ++i) {
            h0->sel[i].above_min = h0->radix[i].above_min = 0xFFFFFFFFu;
            h0->br[i].above_min = h0->br[i].next_min = 0xFFFFFFFFu;
        }
        return b;
    }();
    cudaError_t ce = cudaMemcpyAsync(d, init_block.data(), head, cudaMemcpyHostToDevice, s); // pageable source: staged before the call returns
    const int sms = sm_count(dev);
    StatState h[2];
    StatBracket hb[2];
    memset(hb, 0, sizeof hb);
    auto radix = [&](int which, const StatState *prev) { // which: 0 vmed, 1 vmad; vmin / vmax are on the host by now
        StatParams q{vmin, vmax, which, 0, nullptr};
        // the median of the normalised values is the normalised median pixel when normalisation is monotone and drops nothing but NaN
        q.raw = which == 0 && std::isfinite(vmin) && std::isfinite(vmax) && std::isfinite(vmax - vmin) && vmax > vmin;
        return stat_select_launch(img->ptr, n, q, prev, &hd->radix[which], hd->hist_radix, cand, cand_cap, sms, s);
    };
    if (one_pass) {
        const bool memo = img->range_valid && img->range; // left by the kernel that wrote the pixels: no pass over the image
        const unsigned *mm = memo ? img->range : hd->mm;
        if (ce == cudaSuccess && !memo) ce = minmax_launch(img->ptr, n, hd->mm, sms, s);
        StatParams q{0.0f, 0.0f, 0, 0, mm};
        if (ce == cudaSuccess) ce = stat_bracket_launch(img->ptr, n, q, nullptr, &hd->sel[0], &hd->br[0], hd->hist_br[0], cand, cand_cap, hd->binbuf[0], sms, s);
        q.use_vmed = 1;
        if (ce == cudaSuccess) ce = stat_bracket_launch(img->ptr, n, q, &hd->sel[0], &hd->sel[1], &hd->br[1], hd->hist_br[1], cand, cand_cap, hd->binbuf[1], sms, s);
        unsigned hm[4] = {0, 0, 0, 0};
        // one asynchronous copy into this thread's pinned read-back page (plus the memo's words when the limits came from the
        // image's range memo), one synchronisation
        constexpr size_t kOut = offsetof(Head, mm) + sizeof(hd->mm);
        static_assert(offsetof(Head, sel) == 0 && offsetof(Head, br) == sizeof(StatState) * 2 && offsetof(Head, mm) == offsetof(Head, br) + sizeof(StatBracket) * 2,
                      "sel, br, mm are contiguous at the start of the block");
        static_assert(kOut + 16 <= kReadbackBytes, "fits the read-back page");
        char *rb = nullptr;
        if (ce == cudaSuccess)
            if (Error e = readback_page(&rb)) {
                cudaFreeAsync(d, s);
                return e;
            }
        if (ce == cudaSuccess) ce = cudaMemcpyAsync(rb, d, kOut, cudaMemcpyDeviceToHost, s);
        if (ce == cudaSuccess && memo) ce = cudaMemcpyAsync(rb + kOut, mm, sizeof hm, cudaMemcpyDeviceToHost, s);
        if (ce == cudaSuccess) ce = cudaStreamSynchronize(s);
        if (ce == cudaSuccess) {
            memcpy(h, rb + offsetof(Head, sel), sizeof h);
            memcpy(hb, rb + offsetof(Head, br), sizeof hb);
            memcpy(hm, memo ? rb + kOut : rb + offsetof(Head, mm), sizeof hm);
        }
        // the limits as op_min_max reports them, for the fall-back below
        any = hm[2] != 0;
        vmin = any ? minmax_decode_min(hm[0]) : std::numeric_limits<float>::quiet_NaN();
        vmax = any ? minmax_decode_max(hm[1]) : std::numeric_limits<float>::quiet_NaN();
        // a bracket that missed the wanted rank (about one image in 16000) or could not hold its keys (heavy ties): the radix select
        if (ce == cudaSuccess && hb[0].fallback) {
            ce = radix(0, nullptr);
            if (ce == cudaSuccess) ce = radix(1, &hd->radix[0]);
            if (ce == cudaSuccess) ce = cudaMemcpyAsync(h, hd->radix, sizeof h, cudaMemcpyDeviceToHost, s);
            if (ce == cudaSuccess) ce = cudaStreamSynchronize(s);
        } else if (ce == cudaSuccess && !h[0].error && hb[1].fallback) {
            ce = radix(1, &hd->sel[0]);
            if (ce == cudaSuccess) ce = cudaMemcpyAsync(&h[1], &hd->radix[1], sizeof h[1], cudaMemcpyDeviceToHost, s);
            if (ce == cudaSuccess) ce = cudaStreamSynchronize(s);
        }
    } else {
        if (ce == cudaSuccess) ce = radix(0, nullptr);
        if (ce == cudaSuccess) ce = radix(1, &hd->radix[0]);
        if (ce == cudaSuccess) ce = cudaMemcpyAsync(h, hd->radix, sizeof h, cudaMemcpyDeviceToHost, s);
        if (ce == cudaSuccess) ce = cudaStreamSynchronize(s);
    }
    cudaFreeAsync(d, s);
    if (ce != cudaSuccess) return cuda_error(ce, "stat_select");
    image_note_read(img, s);
    if (h[0].error) return make_error(AFK_ERR_UNEXPECTED_INPUT, "Empty data!, med"); // lims.rs:53
    if (h[1].error) return make_error(AFK_ERR_UNEXPECTED_INPUT, "Empty data!, mad"); // lims.rs:92
    *vmed = h[0].result;
    *vmad = h[1].result;
    return ok();
}

// ---- zscale (zscale.rs:45-242), f32 throughout, sums left to right like Iterator::sum ---------------------------------------
static int rust_f32_as_i32(float f) {
    if (f != f) return 0;
    if (f >= 2147483648.0f) return std::numeric_limits<int>::max();
    if (f <= -2147483648.0f) return std::numeric_limits<int>::min();
    return (int)f;
}
static uint64_t rust_f32_as_usize(float f) {
    if (!(f > 0.0f)) return 0;
    if (f >= 18446744073709551616.0f) return std::numeric_limits<uint64_t>::max();
    return (uint64_t)f;
}

// zsc_compute_sigma (zscale.rs:205-241)
static void zsc_compute_sigma(const std::vector<float> &flat, const std::vector<int> &badpix, int *ngoodpix, bool *has_sigma, float *sigma) {
    float sumz = 0.0f, sumsq = 0.0f;
    int ng = 0;
    for (size_t i = 0; i < flat.size(); ++i)
        if (badpix[i] == 0) {
            sumz = sumz + flat[i];
            ++ng;
        }
    for (size_t i = 0; i < flat.size(); ++i)
        if (badpix[i] == 0) sumsq = sumsq + flat[i] * flat[i];
    *ngoodpix = ng;
    *has_sigma = false;
    if (ng >= 2) {
        // i32 product like the reference: ngoodpix * (ngoodpix - 1)
        const float temp = sumsq / (float)(ng - 1) - sumz * sumz / (float)(ng * (ng - 1));
        *has_sigma = true;
        *sigma = temp < 0.0f ? 0.0f : sqrtf(temp);
    }
}

// zsc_fit_line (zscale.rs:127-203)
static void zsc_fit_line(const std::vector<float> &samples, int npix, float krej, int ngrow, int maxiter, int min_npixels, float max_reject,
                         int *ngoodpix_out, float *zslope_out) {
    const float xscale = 2.0f / (float)(npix - 1);
    std::vector<float> xnorm((size_t)npix);
    for (int v = 0; v < npix; ++v) xnorm[(size_t)v] = (float)v * xscale - 1.0f;
    int ngoodpix = npix;
    const int minpix = std::max(min_npixels, rust_f32_as_i32((float)npix * max_reject));
    const int last_ngoodpix = npix + 1;
    std::vector<int> badpix((size_t)npix, 0);
    float slope = 0.0f;
    std::vector<float> flat((size_t)npix);
    for (int it = 0; it < maxiter; ++it) {
        if (ngoodpix >= last_ngoodpix || ngoodpix < minpix) break;
        int sum = 0;
        float sumx = 0.0f, sumxx = 0.0f, sumxy = 0.0f, sumy = 0.0f;
        for (int i = 0; i < npix; ++i)
            if (badpix[(size_t)i] == 0) {
                ++sum;
                sumx = sumx + xnorm[(size_t)i];
            }
        for (int i = 0; i < npix; ++i)
            if (badpix[(size_t)i] == 0) sumxx = sumxx + xnorm[(size_t)i] * xnorm[(size_t)i];
        for (int i = 0; i < npix; ++i)
            if (badpix[(size_t)i] == 0) sumxy = sumxy + xnorm[(size_t)i] * samples[(size_t)i];
        for (int i = 0; i < npix; ++i)
            if (badpix[(size_t)i] == 0) sumy = sumy + samples[(size_t)i];
        const float delta = (float)sum * sumxx - sumx * sumx;
        const float intercept = (sumxx * sumy - sumx * sumxy) / delta;
        slope = ((float)sum * sumxy - sumx * sumy) / delta;
        for (int i = 0; i < npix; ++i) flat[(size_t)i] = samples[(size_t)i] - (xnorm[(size_t)i] * slope + intercept);
        int ng;
        bool has_sigma;
        float sigma = 0.0f;
        zsc_compute_sigma(flat, badpix, &ng, &has_sigma, &sigma);
        ngoodpix = ng;
        if (has_sigma) { // (Some(mean), Some(sigma))
            const float threshold = sigma * krej, lcut = -threshold, hcut = threshold;
            std::vector<int> next((size_t)npix);
            for (int i = 0; i < npix; ++i) next[(size_t)i] = (flat[(size_t)i] < lcut || flat[(size_t)i] > hcut) ? 1 : 0;
            // convolve(next_badpix, [1, ngrow]): ret[head] = target[head] + target[head-1] * ngrow
            for (int head = 0; head < npix; ++head) {
                int data = next[(size_t)head];
                if (head >= 1) data += next[(size_t)head - 1] * ngrow;
                badpix[(size_t)head] = data;
            }
            ngoodpix = 0;
            for (int i = 0; i < npix; ++i) ngoodpix += badpix[(size_t)i] == 0;
        }
    }
    *ngoodpix_out = ngoodpix;
    *zslope_out = slope * xscale;
}

static Error zscale_run(const afk_image *img, int nsamples, float contrast, int *has, float *z1, float *z2) {
    *has = 0;
    if (img->ndim != 2) return make_error(AFK_ERR_PANIC, "called `Result::unwrap()` on an `Err` value: Msg(\"zsc sample failed.\")");
    const int64_t nc = img->shape[0], nl = img->shape[1];
    // stride = 1.max((((nc - 1) * (nl - 1)) as f32 / maxpix as f32).sqrt() as usize)   (usize arithmetic wraps for empty axes:
    // an empty image has no samples either way)
    std::vector<float> sample;
    if (nc > 0 && nl > 0) {
        const uint64_t raw = rust_f32_as_usize(sqrtf((float)((uint64_t)(nc - 1) * (uint64_t)(nl - 1)) / (float)nsamples));
        const int64_t stride = (int64_t)std::min<uint64_t>(std::max<uint64_t>(1, raw), (uint64_t)std::numeric_limits<int64_t>::max());
        const int64_t rows_s = (nc - 1) / stride + 1, cols_s = (nl - 1) / stride + 1;
        std::vector<float> grid((size_t)(rows_s * cols_s));
        cudaStream_t s;
        int dev;
        if (Error e = stats_ctx(img, &s, &dev)) return e;
        cudaError_t ce;
        if (grid.size() * sizeof(float) <= kReadbackBytes - 1024) {
            // the usual ~1000 samples: the kernel stores them straight into this thread's mapped read-back page (one launch,
            // one synchronisation; no allocation or copy)
            char *page = nullptr;
            if (Error e = readback_page(&page)) return e;
            ce = zsample_launch(img->ptr, nc, nl, stride, rows_s, cols_s, reinterpret_cast<float *>(page + 1024), s);
            if (ce == cudaSuccess) ce = cudaStreamSynchronize(s);
            if (ce == cudaSuccess) memcpy(grid.data(), page + 1024, grid.size() * sizeof(float));
        } else {
            float *d = nullptr;
            AFK_CUDA(cudaMallocAsync(reinterpret_cast<void **>(&d), grid.size() * sizeof(float), s));
            ce = zsample_launch(img->ptr, nc, nl, stride, rows_s, cols_s, d, s);
            if (ce == cudaSuccess) ce = cudaMemcpyAsync(grid.data(), d, grid.size() * sizeof(float), cudaMemcpyDeviceToHost, s);
            if (ce == cudaSuccess) ce = cudaStreamSynchronize(s);
            cudaFreeAsync(d, s);
        }
        if (ce != cudaSuccess) return cuda_error(ce, "zsample_kernel");
        image_note_read(img, s);
        for (float v : grid)
            if (v == v) sample.push_back(v);
    }
    // v.truncate(maxpix as usize): a negative i32 becomes a huge usize and keeps everything
    if (nsamples >= 0 && sample.size() > (size_t)nsamples) sample.resize((size_t)nsamples);
    std::sort(sample.begin(), sample.end()); // no NaN left: partial_cmp never fails
    const size_t npix = sample.size();
    if (npix == 0) return ok(); // (None, None)
    const float zmin = sample.front(), zmax = sample.back();
    const size_t center_pixel = (npix - 1) / 2;
    float median;
    if (npix % 2 == 1) median = sample[center_pixel];
    else median = 0.5f * (sample[center_pixel] + sample[center_pixel + 1]);
    const int MIN_NPIXELS = 5, MAX_ITERATIONS = 5;
    const float MAX_REJECT = 0.5f, KREJ = 2.5f;
    const int minpix = std::max(MIN_NPIXELS, rust_f32_as_i32((float)npix * MAX_REJECT));
    const int ngrow = std::max(1, rust_f32_as_i32((float)npix * 0.01f));
    int ngoodpix;
    float zslope;
    zsc_fit_line(sample, (int)npix, KREJ, ngrow, MAX_ITERATIONS, MIN_NPIXELS, MAX_REJECT, &ngoodpix, &zslope);
    if (ngoodpix < minpix) {
        *z1 = zmin;
        *z2 = zmax;
    } else {
        if (contrast > 0.0f) zslope = zslope / contrast;
        // (center_pixel - 1) is usize arithmetic: it would overflow (panic in debug) for npix <= 2, which cannot pass ngoodpix >= 5
        *z1 = fmaxf(zmin, median - (float)(center_pixel - 1) * zslope);
        *z2 = fminf(zmax, median + (float)(npix - center_pixel) * zslope);
    }
    *has = 1;
    return ok();
}

} // namespace afk

using namespace afk;

extern "C" {

afk_status afk_image_histogram(const afk_image *img, float vmin, float vmax, uint64_t counts[100], float starts[100], float ends[100]) {
    if (!img || !counts) return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "null argument"));
    memset(counts, 0, 100 * sizeof(uint64_t));
    return fail(histogram_run(img, 1, 100, vmin, vmax, counts, starts, ends));
}

afk_status afk_image_histogram_color(const afk_image *img, float vmin, float vmax, uint64_t counts[768], float starts[256], float ends[256]) {
    if (!img || !counts) return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "null argument"));
    if (img->ndim != 3 || img->shape[0] != 3) return fail(make_error(AFK_ERR_SHAPE, "histogram_color needs a [3, ny, nx] image"));
    memset(counts, 0, 768 * sizeof(uint64_t));
    return fail(histogram_run(img, 3, 256, vmin, vmax, counts, starts, ends));
}

afk_status afk_image_vmed_vmad(const afk_image *img, float *vmed, float *vmad) {
    if (!img || !vmed || !vmad) return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "null argument"));
    return fail(vmed_vmad_run(img, vmed, vmad));
}

afk_status afk_image_zscale(const afk_image *img, int nsamples, float contrast, int *has_limits, float *z1, float *z2) {
    if (!img || !has_limits || !z1 || !z2) return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "null argument"));
    return fail(zscale_run(img, nsamples, contrast, has_limits, z1, z2));
}

} // extern "C"
