// display.cpp -- the aflak_plot colour-mapping path: ColorLUT (host-built 65536-entry table, device copy) and the
// fused make_raw_image kernel launch.  Headless: the result is a device (or host) byte buffer, not a GL texture.
// Reference citations are relative to /root/reference/src/.
#include <cmath>
#include <cstring>
#include <limits>

#include <array>
#include <algorithm>
#include <vector>
#include "host.h"

namespace afk {

constexpr int kLutSize = 65536; // LUT_SIZE, lut.rs:6

struct Stop {
    float v;
    uint8_t c[3];
};

} // namespace afk

struct afk_lut {
    std::vector<afk::Stop> gradient;
    std::vector<uint8_t> table; // 65536 x 3
    float lims[3] = {0.0f, 0.5f, 1.0f};
    // device copies (packed r | g<<8 | b<<16 | 0xFF<<24), one per device, built lazily
    std::mutex mu;
    uint32_t *dev_table[16] = {};
    bool dev_valid[16] = {};
};

namespace afk {

// Rust `f as u8`: truncate, saturate, NaN -> 0
static inline uint8_t as_u8(float f) {
    if (!(f == f) || f <= 0.0f) return 0;
    if (f >= 255.0f) return 255;
    return (uint8_t)f;
}

// StopIter (lut.rs:251-286)
static bool stop_iter(const afk_lut *l, size_t i, float *v, const uint8_t **c) {
    const std::vector<Stop> &g = l->gradient;
    if (g.empty()) return false;
    if (i == 0) {
        *v = 0.0f;
        *c = g[0].c;
        return true;
    }
    if (i - 1 == g.size()) {
        *v = 1.0f;
        *c = g.back().c;
        return true;
    }
    if (i - 1 < g.size()) {
        float span = l->lims[2] - l->lims[0];
        float t = span * g[i - 1].v;
        *v = l->lims[0] + t;
        *c = g[i - 1].c;
        return true;
    }
    return false;
}

// color_at_init (lut.rs:162-188)
static void color_at_init(const afk_lut *l, float point, uint8_t out[3]) {
    for (size_t i = 0;; ++i) {
        float v1, v2;
        const uint8_t *c1, *c2;
        if (!stop_iter(l, i, &v1, &c1) || !stop_iter(l, i + 1, &v2, &c2)) break;
        float dv = v2 - v1;
        if (v1 <= point && point <= v2) {
            if (dv == 0.0f) {
                memcpy(out, c1, 3);
            } else {
                float dp = point - v1;
                float coef = dp / dv;
                for (int ch = 0; ch < 3; ++ch) {
                    float a = (float)c1[ch], b = (float)c2[ch];
                    float d = b - a;
                    d = d * coef;
                    out[ch] = as_u8(a + d);
                }
            }
            return;
        }
    }
    out[0] = out[1] = out[2] = 0;
}

// lut_init (lut.rs:190-194)
static void lut_init(afk_lut *l) {
    l->table.resize((size_t)kLutSize * 3);
    for (int i = 0; i < kLutSize; ++i) color_at_init(l, (float)i / (float)(kLutSize - 1), &l->table[(size_t)i * 3]);
    std::lock_guard<std::mutex> lk(l->mu);
    for (int d = 0; d < 16; ++d) l->dev_valid[d] = false;
}

static float mtf_host(float m, float x) { // lut.rs:152-160
    if (x >= 1.0f) return 1.0f;
    if (x <= 0.0f) return 0.0f;
    float num = (m - 1.0f) * x;
    float den = 2.0f * m;
    den = den - 1.0f;
    den = den * x;
    den = den - m;
    return num / den;
}

static uint32_t lut_index_host(const afk_lut *l, float point) { // color_at, lut.rs:137-149
    float c = (point - l->lims[0]) / (l->lims[2] - l->lims[0]);
    float i = mtf_host(l->lims[1], c) * (float)(kLutSize - 1);
    if (i < 0.0f) i = 0.0f;
    uint32_t idx;
    if (!(i > 0.0f)) idx = 0;
    else if (i >= 65535.0f) idx = 65535;
    else idx = (uint32_t)i;
    return idx;
}

static Error lut_device(afk_lut *l, int dev, cudaStream_t s, const uint32_t **out) {
    std::lock_guard<std::mutex> lk(l->mu);
    if (!l->dev_table[dev]) AFK_CUDA(cudaMalloc(reinterpret_cast<void **>(&l->dev_table[dev]), (size_t)kLutSize * 4));
    if (!l->dev_valid[dev]) {
        std::vector<uint32_t> packed(kLutSize);
        for (int i = 0; i < kLutSize; ++i)
            packed[i] = (uint32_t)l->table[3 * i] | ((uint32_t)l->table[3 * i + 1] << 8) | ((uint32_t)l->table[3 * i + 2] << 16) | 0xFF000000u;
        AFK_CUDA(cudaMemcpyAsync(l->dev_table[dev], packed.data(), (size_t)kLutSize * 4, cudaMemcpyHostToDevice, s));
        AFK_CUDA(cudaStreamSynchronize(s)); // `packed` is a stack-lifetime staging buffer
        l->dev_valid[dev] = true;
    }
    *out = l->dev_table[dev];
    return ok();
}

struct Builtin {
    const char *name;
    std::vector<Stop> stops;
};

// BuiltinLUT gradients (lut.rs:61-113) and names (lut.rs:46-59)
static const std::vector<Builtin> &builtins() {
    static const std::vector<Builtin> b = {
        {"Grey", {{0.0f, {0, 0, 0}}, {1.0f, {255, 255, 255}}}},
        {"GreyClip", {{0.0f, {0, 0, 0}}, {0.99f, {255, 255, 255}}, {1.0f, {255, 0, 0}}}},
        {"Thermal", {{0.0f, {0, 0, 0}}, {1.0f / 3.0f, {185, 0, 0}}, {2.0f / 3.0f, {255, 220, 0}}, {1.0f, {255, 255, 255}}}},
        {"Flame", {{0.0f, {0, 0, 0}}, {0.2f, {7, 0, 220}}, {0.5f, {236, 0, 134}}, {0.8f, {246, 246, 0}}, {1.0f, {255, 255, 255}}}},
        {"Yellowy", {{0.0f, {0, 0, 0}}, {0.25f, {32, 0, 129}}, {0.5f, {115, 15, 255}}, {0.75f, {255, 255, 0}}, {1.0f, {255, 255, 255}}}},
        {"HeatMap", {{0.0f, {1, 1, 85}}, {0.1f, {0, 0, 255}}, {0.25f, {0, 255, 255}}, {0.5f, {0, 255, 0}}, {0.75f, {255, 255, 0}},
                     {0.9f, {255, 0, 0}}, {0.99f, {108, 6, 10}}, {1.0f, {255, 255, 255}}}},
        {"HeatMap_Inv", {{0.0f, {108, 6, 10}}, {0.1f, {255, 0, 0}}, {0.25f, {255, 255, 0}}, {0.5f, {0, 255, 0}}, {0.75f, {0, 255, 255}},
                         {0.89f, {0, 0, 255}}, {0.99f, {1, 1, 85}}, {1.0f, {255, 255, 255}}}},
        {"Red", {{0.0f, {0, 0, 0}}, {1.0f, {255, 0, 0}}}},
        {"Blue", {{0.0f, {0, 0, 0}}, {1.0f, {0, 0, 255}}}},
        {"Green", {{0.0f, {0, 0, 0}}, {1.0f, {0, 255, 0}}}},
    };
    return b;
}

static Error colormap_run(const afk_image *img, const afk_lut *const *luts, int nlut, const float *vmin, const float *vmax, int flip_y,
                          int channels, const afk_colormap_chain *chain, uint8_t *out_device, cudaStream_t s, int dev) {
    if (channels != 3 && channels != 4) return make_error(AFK_ERR_UNEXPECTED_INPUT, "channels must be 3 (RGB8) or 4 (RGBA8)");
    const bool rgb3 = nlut == 3;
    if (!rgb3 && img->ndim != 2) return make_error(AFK_ERR_UNEXPECTED_INPUT, fmt("'image' is a %d-dimensional image, while dimension 2 was expected", img->ndim));
    if (rgb3 && (img->ndim != 3 || img->shape[0] != 3)) return make_error(AFK_ERR_UNEXPECTED_INPUT, "colour image must have shape [3, ny, nx]");
    if (img->device != dev) return make_error(AFK_ERR_CUDA, "CUDA: image lives on another device");
    ColormapArgs a{};
    a.img = img->ptr;
    a.ny = img->shape[img->ndim - 2];
    a.nx = img->shape[img->ndim - 1];
    uint32_t *lut3 = nullptr;
    Error err = ok();
    float *dtone = nullptr;
    do {
        if (rgb3) {
            // three tables side by side
            cudaError_t ce = cudaMallocAsync(reinterpret_cast<void **>(&lut3), (size_t)3 * kLutSize * 4, s);
            if (ce != cudaSuccess) { err = cuda_error(ce, "cudaMallocAsync(lut3)"); break; }
            for (int c = 0; c < 3 && !err; ++c) {
                const uint32_t *t = nullptr;
                if ((err = lut_device(const_cast<afk_lut *>(luts[c]), dev, s, &t))) break;
                ce = cudaMemcpyAsync(lut3 + (size_t)c * kLutSize, t, (size_t)kLutSize * 4, cudaMemcpyDeviceToDevice, s);
                if (ce != cudaSuccess) err = cuda_error(ce, "cudaMemcpyAsync(lut3)");
            }
            if (err) break;
            a.lut = lut3;
        } else {
            if ((err = lut_device(const_cast<afk_lut *>(luts[0]), dev, s, &a.lut))) break;
        }
        for (int c = 0; c < 3; ++c) {
            const afk_lut *l = luts[rgb3 ? c : 0];
            a.vmin[c] = vmin[rgb3 ? c : 0];
            a.vmax[c] = vmax[rgb3 ? c : 0];
            for (int k = 0; k < 3; ++k) a.lims[c][k] = l->lims[k];
        }
        a.rgb3 = rgb3;
        a.flip_y = flip_y;
        a.channels = channels;
        a.out = out_device;
        if (chain) {
            if (!chain->tone_table || chain->tone_len < 1 || chain->tone_len > (size_t)kMaxToneTable) {
                err = make_error(AFK_ERR_UNEXPECTED_INPUT, "bad tone table in colormap chain");
                break;
            }
            // The table is a GUI value that changes when the user drags a control point, not per frame: keep the last one
            // uploaded on this stream and compare (1 KiB memcmp) instead of an allocation + upload + free per call.
            StreamScratch &sc = tls().scratch[std::make_pair(dev, s)];
            const bool same = sc.tone_dev && sc.tone_host.size() == chain->tone_len &&
                              memcmp(sc.tone_host.data(), chain->tone_table, chain->tone_len * sizeof(float)) == 0;
            if (!same) {
                cudaError_t ce = cudaSuccess;
                sc.tone_host.clear(); // not valid until the copy below has been enqueued
                if (sc.tone_cap < chain->tone_len) {
                    if (sc.tone_dev) cudaFreeAsync(sc.tone_dev, s);
                    sc.tone_dev = nullptr;
                    sc.tone_cap = 0;
                    const size_t cap = chain->tone_len < 1024 ? 1024 : chain->tone_len;
                    ce = cudaMallocAsync(reinterpret_cast<void **>(&sc.tone_dev), cap * sizeof(float), s);
                    if (ce == cudaSuccess) sc.tone_cap = cap;
                    else sc.tone_dev = nullptr;
                }
                // pageable source: the call returns once the entries are staged, and kernels enqueued earlier on `s` that
                // read the previous table finish before the copy runs (stream order)
                if (ce == cudaSuccess) ce = cudaMemcpyAsync(sc.tone_dev, chain->tone_table, chain->tone_len * sizeof(float), cudaMemcpyHostToDevice, s);
                if (ce != cudaSuccess) { err = cuda_error(ce, "tone table upload"); break; }
                sc.tone_host.assign(chain->tone_table, chain->tone_table + chain->tone_len);
            }
            dtone = sc.tone_dev;
            a.chain = 1;
            if (chain->logscale_vmin != chain->logscale_vmin && chain->logscale_vmax != chain->logscale_vmax) {
                // NaN, NaN: the chain starts at image_min_max (src/templates.rs: min/max feed convert_to_logscale).  Its result stays
                // on the device -- no read-back between the two kernels.
                if ((err = image_wait(img, s))) break;
                if (img->range_valid && img->range) {
                    a.ls_minmax = img->range; // range memo left by the kernel that wrote the pixels
                } else {
                    if (!sc.minmax_words) {
                        cudaError_t ce = cudaMallocAsync(reinterpret_cast<void **>(&sc.minmax_words), 4 * sizeof(unsigned), s);
                        if (ce != cudaSuccess) { sc.minmax_words = nullptr; err = cuda_error(ce, "cudaMallocAsync(min/max words)"); break; }
                    }
                    cudaError_t ce = cudaMemsetAsync(sc.minmax_words, 0, 4 * sizeof(unsigned), s);
                    if (ce == cudaSuccess) ce = minmax_launch(img->ptr, img->len(), sc.minmax_words, sm_count(dev), s);
                    if (ce != cudaSuccess) { err = cuda_error(ce, "minmax_kernel launch"); break; }
                    a.ls_minmax = sc.minmax_words;
                }
            }
            a.ls_a = chain->logscale_a; a.ls_min = chain->logscale_vmin; a.ls_max = chain->logscale_vmax;
            a.nm_min = chain->normalize_min; a.nm_max = chain->normalize_max;
            a.tone_table = dtone;
            a.tone_len = (int)chain->tone_len;
            // apply_tone_curve indexes its table unchecked (lib.rs:1866): an input above the table panics in the reference.  The
            // kernel records that in this thread's panic word (mapped host memory, written only when it happens); it is reported
            // at the next host synchronisation (afk_synchronize, afk_colormap_host) -- no read-back on the steady-state path.
            if ((err = panic_word_get(dev, &a.oob_flag))) break;
            tls().panic_tone_len = chain->tone_len;
        }
        if ((err = image_wait(img, s))) break;
        cudaError_t ce = colormap_launch(a, sm_count(dev), s);
        if (ce != cudaSuccess) { err = cuda_error(ce, "colormap_kernel launch"); break; }
        image_note_read(img, s);
    } while (0);
    if (lut3) cudaFreeAsync(lut3, s);
    return err;
}

static Error resolve_limits(const afk_image *img, float *vmin, float *vmax) {
    // Image::new (image.rs:143-194): vmin/vmax from lims::get_vmin/get_vmax when not supplied
    if (*vmin == *vmin && *vmax == *vmax) return ok();
    float mn, mx;
    bool any;
    if (Error e = op_min_max(img, &mn, &mx, &any)) return e;
    if (*vmin != *vmin) *vmin = mn;
    if (*vmax != *vmax) *vmax = mx;
    if (*vmin != *vmin || *vmax != *vmax) return make_error(AFK_ERR_UNEXPECTED_INPUT, "vmin, vmax not set"); // image.rs:46-48
    return ok();
}

} // namespace afk

using namespace afk;

extern "C" {

int afk_set_display_exact(int on) {
    set_display_exact(on != 0);
    return AFK_OK;
}
int afk_get_display_exact(void) { return display_exact() ? 1 : 0; }

afk_lut *afk_lut_linear(const float *stop_values, const uint8_t *stop_rgb, size_t n_stops) {
    afk_lut *l = new afk_lut();
    for (size_t i = 0; i < n_stops; ++i) {
        Stop s;
        s.v = stop_values[i];
        memcpy(s.c, stop_rgb + 3 * i, 3);
        l->gradient.push_back(s);
    }
    lut_init(l);
    return l;
}

afk_lut *afk_lut_builtin(const char *name) {
    if (!name) return nullptr;
    for (const Builtin &b : builtins())
        if (strcmp(b.name, name) == 0) {
            afk_lut *l = new afk_lut();
            l->gradient = b.stops;
            lut_init(l);
            return l;
        }
    fail(make_error(AFK_ERR_UNEXPECTED_INPUT, fmt("unknown builtin LUT '%s'", name)));
    return nullptr;
}

size_t afk_lut_builtin_count(void) { return builtins().size(); }
const char *afk_lut_builtin_name(size_t i) { return i < builtins().size() ? builtins()[i].name : nullptr; }

void afk_lut_release(afk_lut *l) {
    if (!l) return;
    for (int d = 0; d < 16; ++d)
        if (l->dev_table[d]) cudaFree(l->dev_table[d]);
    delete l;
}

static float clamp01(float v) { return v < 0.0f ? 0.0f : (v > 1.0f ? 1.0f : v); }

void afk_lut_set_lims(afk_lut *l, float mn, float mid, float mx) {
    // set_min / set_mid / set_max (lut.rs:202-236)
    if (!l) return;
    mn = clamp01(mn);
    if (mn > l->lims[2]) l->lims[2] = mn;
    l->lims[0] = mn;
    l->lims[1] = clamp01(mid);
    mx = clamp01(mx);
    if (mx < l->lims[0]) l->lims[0] = mx;
    l->lims[2] = mx;
}

void afk_lut_lims(const afk_lut *l, float out[3]) {
    if (l) memcpy(out, l->lims, sizeof l->lims);
}

int afk_lut_set_gradient(afk_lut *l, const float *stop_values, const uint8_t *stop_rgb, size_t n_stops) {
    // set_gradient (lut.rs:242-245): the table is rebuilt with the lims in force at this moment
    if (!l) return AFK_ERR_UNEXPECTED_INPUT;
    l->gradient.clear();
    for (size_t i = 0; i < n_stops; ++i) {
        Stop s;
        s.v = stop_values[i];
        memcpy(s.c, stop_rgb + 3 * i, 3);
        l->gradient.push_back(s);
    }
    lut_init(l);
    return AFK_OK;
}

void afk_lut_color_at(const afk_lut *l, float point, uint8_t rgb[3]) {
    uint32_t i = lut_index_host(l, point);
    memcpy(rgb, &l->table[(size_t)i * 3], 3);
}

void afk_lut_color_at_bounds(const afk_lut *l, float point, float vmin, float vmax, uint8_t rgb[3]) {
    float p = point; // util::clamp (util.rs:3-14)
    if (p < vmin) p = vmin;
    else if (p > vmax) p = vmax;
    afk_lut_color_at(l, (p - vmin) / (vmax - vmin), rgb);
}

void afk_lut_table(const afk_lut *l, uint8_t *out_rgb) { memcpy(out_rgb, l->table.data(), l->table.size()); }

size_t afk_lut_bounds(const afk_lut *l, float *v1, uint8_t *c1, float *v2, uint8_t *c2, size_t cap) {
    size_t n = 0;
    for (size_t i = 0;; ++i) {
        float a, b;
        const uint8_t *ca, *cb;
        if (!stop_iter(l, i, &a, &ca) || !stop_iter(l, i + 1, &b, &cb)) break;
        if (n < cap) {
            v1[n] = a;
            v2[n] = b;
            memcpy(c1 + 3 * n, ca, 3);
            memcpy(c2 + 3 * n, cb, 3);
        }
        ++n;
    }
    return n;
}

// UiToneCurve::create_curve (imgui_tone_curve/src/lib.rs:68-158): the table `arr` of a ToneCurveState from its control
// points.  GUI-side code in the reference (the widget rebuilds `arr` whenever a point moves); offered here so that a
// headless host can build the value apply_tone_curve consumes.  Every step is one f32 operation, in the source's order.
static bool tone_sorted_points(const float *cp_xy, size_t n, std::vector<std::array<float, 2>> &out) {
    out.resize(n);
    for (size_t i = 0; i < n; ++i) {
        out[i] = {cp_xy[2 * i], cp_xy[2 * i + 1]};
        if (out[i][0] != out[i][0]) return false; // partial_cmp(..).unwrap() panics on NaN
    }
    std::stable_sort(out.begin(), out.end(), [](const std::array<float, 2> &a, const std::array<float, 2> &b) { return a[0] < b[0]; });
    return true;
}

static size_t rust_as_usize(float v) { // `as usize`: saturating, NaN -> 0
    if (!(v > 0.0f)) return 0;
    if (v >= 18446744073709551616.0f) return SIZE_MAX;
    return (size_t)v;
}

afk_status afk_tone_curve_from_points(const float *cp_xy, size_t n_cp, float *arr_out, size_t cap, size_t *n_out) {
    if ((!cp_xy && n_cp) || !n_out) return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "null argument to afk_tone_curve_from_points"));
    std::vector<float> ret;
    std::vector<std::array<float, 2>> cp;
    if (n_cp == 2) {
        const float x0 = cp_xy[0], y0 = cp_xy[1], x1 = cp_xy[2], y1 = cp_xy[3];
        const float slope = (y1 - y0) / (x1 - x0);
        for (int i = 0; i < 256; ++i) {
            const float t = (float)i / 256.0f;
            ret.push_back((t * slope + y0) * 256.0f);
        }
    } else if (n_cp == 3) {
        if (!tone_sorted_points(cp_xy, n_cp, cp)) return fail(make_error(AFK_ERR_PANIC, "called `Option::unwrap()` on a `None` value"));
        const float x1 = cp[0][0], x2 = cp[1][0], x3 = cp[2][0], y1 = cp[0][1], y2 = cp[1][1], y3 = cp[2][1];
        const float a = ((y1 - y2) * (x1 - x3) - (y1 - y3) * (x1 - x2)) / ((x1 - x2) * (x1 - x3) * (x2 - x3));
        const float b = (y1 - y2) / (x1 - x2) - a * (x1 + x2);
        const float c = y1 - a * x1 * x1 - b * x1;
        for (int i = 0; i < 256; ++i) {
            const float t = (float)i / 256.0f;
            ret.push_back((a * t * t + b * t + c) * 256.0f);
        }
    } else if (n_cp > 3) {
        if (!tone_sorted_points(cp_xy, n_cp, cp)) return fail(make_error(AFK_ERR_PANIC, "called `Option::unwrap()` on a `None` value"));
        for (size_t counter = 0; counter + 1 < n_cp; ++counter) {
            if (counter == 0) {
                const auto p0 = cp[0], p1 = cp[1], p2 = cp[2];
                const size_t size = rust_as_usize(std::ceil(p1[0] * 256.0f));
                const float b = p0[1] - 2.0f * p1[1] + p2[1];
                const float c = -3.0f * p0[1] + 4.0f * p1[1] - p2[1];
                const float d = 2.0f * p0[1];
                for (size_t i = 0; i < size; ++i) {
                    const float t = (float)i / (float)(size - 1);
                    ret.push_back(((b * t * t) + (c * t) + d) * 128.0f);
                }
            } else if (counter == n_cp - 2) {
                const auto p0 = cp[counter - 1], p1 = cp[counter], p2 = cp[counter + 1];
                const size_t size = rust_as_usize(std::floor(p2[0] * 256.0f) - std::ceil(p1[0] * 256.0f));
                const float b = p0[1] - 2.0f * p1[1] + p2[1];
                const float c = -p0[1] + p2[1];
                const float d = 2.0f * p1[1];
                for (size_t i = 0; i < size; ++i) {
                    const float t = (float)i / (float)(size - 1);
                    ret.push_back(((b * t * t) + (c * t) + d) * 128.0f);
                }
            } else {
                const auto p0 = cp[counter - 1], p1 = cp[counter], p2 = cp[counter + 1], p3 = cp[counter + 2];
                const size_t size = rust_as_usize(std::floor(p2[0] * 256.0f) - std::ceil(p1[0] * 256.0f)) + 1;
                const float a = -p0[1] + 3.0f * p1[1] - 3.0f * p2[1] + p3[1];
                const float b = 2.0f * p0[1] - 5.0f * p1[1] + 4.0f * p2[1] - p3[1];
                const float c = -p0[1] + p2[1];
                const float d = 2.0f * p1[1];
                for (size_t i = 0; i < size; ++i) {
                    const float t = (float)i / (float)(size - 1);
                    ret.push_back(((a * t * t * t) + (b * t * t) + (c * t) + d) * 128.0f);
                }
            }
            if (ret.size() > ((size_t)1 << 24)) return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "tone curve table larger than 2^24 entries"));
        }
    } // 0 or 1 control points: an empty table
    *n_out = ret.size();
    if (arr_out) {
        if (cap < ret.size()) return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, fmt("tone curve table needs %zu entries, the buffer holds %zu", ret.size(), cap)));
        if (!ret.empty()) memcpy(arr_out, ret.data(), ret.size() * sizeof(float));
    }
    return AFK_OK;
}

afk_status afk_image_vmin_vmax(const afk_image *img, float *vmin, float *vmax) {
    if (!img || !vmin || !vmax) return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "null argument"));
    if (img->len() == 0) return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "Could not get vmin")); // lims.rs:15-19
    bool any;
    if (Error e = op_min_max(img, vmin, vmax, &any)) return fail(e);
    return AFK_OK; // NaN/NaN when every pixel is NaN (float_compare_nan_*, lims.rs:108-124)
}

afk_status afk_colormap(const afk_image *img, const afk_lut *lut, float vmin, float vmax, int flip_y, int channels,
                        const afk_colormap_chain *chain, uint8_t *out_device) {
    if (!img || !lut || !out_device) return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "null argument"));
    cudaStream_t s;
    int dev;
    if (Error e = current_stream(&s, &dev)) return fail(e);
    if (!chain)
        if (Error e = resolve_limits(img, &vmin, &vmax)) return fail(e);
    if (chain && (vmin != vmin || vmax != vmax)) return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "vmin, vmax not set"));
    const afk_lut *luts[1] = {lut};
    return fail(colormap_run(img, luts, 1, &vmin, &vmax, flip_y, channels, chain, out_device, s, dev));
}

afk_status afk_colormap_host(const afk_image *img, const afk_lut *lut, float vmin, float vmax, int flip_y, int channels,
                             const afk_colormap_chain *chain, uint8_t *out_host) {
    if (!img || !lut || !out_host) return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "null argument"));
    cudaStream_t s;
    int dev;
    if (Error e = current_stream(&s, &dev)) return fail(e);
    const size_t bytes = (size_t)img->len() * (size_t)channels;
    uint8_t *d = nullptr;
    cudaError_t ce = cudaMallocAsync(reinterpret_cast<void **>(&d), bytes ? bytes : 16, s);
    if (ce != cudaSuccess) return fail(cuda_error(ce, "cudaMallocAsync(rgba)"));
    afk_status st = afk_colormap(img, lut, vmin, vmax, flip_y, channels, chain, d);
    if (st == AFK_OK && bytes) {
        ce = cudaMemcpyAsync(out_host, d, bytes, cudaMemcpyDeviceToHost, s);
        if (ce == cudaSuccess) ce = cudaStreamSynchronize(s);
        if (ce != cudaSuccess) st = fail(cuda_error(ce, "cudaMemcpyAsync(D2H rgba)"));
    }
    if (st == AFK_OK && chain) st = fail(panic_word_take(dev)); // the stream is idle: the tone-table check of the chain
    cudaFreeAsync(d, s);
    return st;
}

afk_status afk_colormap_rgb_host(const afk_image *img, const afk_lut *const lut[3], const float vmin[3], const float vmax[3], int flip_y,
                                 int channels, uint8_t *out_host) {
    if (!img || !lut || !lut[0] || !lut[1] || !lut[2] || !out_host) return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "null argument"));
    cudaStream_t s;
    int dev;
    if (Error e = current_stream(&s, &dev)) return fail(e);
    if (img->ndim != 3) return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "colour image must have shape [3, ny, nx]"));
    const size_t bytes = (size_t)img->shape[1] * (size_t)img->shape[2] * (size_t)channels;
    uint8_t *d = nullptr;
    cudaError_t ce = cudaMallocAsync(reinterpret_cast<void **>(&d), bytes ? bytes : 16, s);
    if (ce != cudaSuccess) return fail(cuda_error(ce, "cudaMallocAsync(rgb)"));
    afk_status st = fail(colormap_run(img, lut, 3, vmin, vmax, flip_y, channels, nullptr, d, s, dev));
    if (st == AFK_OK && bytes) {
        ce = cudaMemcpyAsync(out_host, d, bytes, cudaMemcpyDeviceToHost, s);
        if (ce == cudaSuccess) ce = cudaStreamSynchronize(s);
        if (ce != cudaSuccess) st = fail(cuda_error(ce, "cudaMemcpyAsync(D2H rgb)"));
    }
    cudaFreeAsync(d, s);
    return st;
}

} // extern "C"
