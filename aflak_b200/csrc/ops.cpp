// ops.cpp -- the operator library: host-side mirror of aflak_primitives' TRANSFORMATIONS for the hot path.
// Each operator does the reference's precondition checks (same messages as precond.rs), launches the sm_100a
// kernel on the calling thread's stream and applies the reference's metadata rules.  No CPU fallback.
// Reference citations are relative to /root/reference/src/.
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <limits>

#include "host.h"

namespace afk {

// =============================================================================================================
// Unit algebra: DerivedUnit (unit.rs:39-163) and Unit::{mul,div} (unit.rs:454-495)
// =============================================================================================================
struct Derived {
    bool has_exp = false;
    float pre = 1.0f;
    long p = 0;
    std::vector<std::pair<std::string, long>> derived; // insertion ordered (the reference's HashMap order is unspecified)
    void insert(const std::string &k, long v) {
        for (auto &kv : derived)
            if (kv.first == k) {
                kv.second = v; // HashMap::insert overwrites
                return;
            }
        derived.emplace_back(k, v);
    }
    long *find(const std::string &k) {
        for (auto &kv : derived)
            if (kv.first == k) return &kv.second;
        return nullptr;
    }
};

static bool is_alpha_or_bar(char c) { return (c >= 'A' && c <= 'Z') || (c >= 'a' && c <= 'z') || c == '|'; }

static Derived derived_parse(const std::string &s) {
    Derived d;
    std::string exp, dunit = s;
    size_t sp = s.find(' ');
    bool has_exp_part = sp != std::string::npos;
    if (has_exp_part) {
        exp = s.substr(0, sp);
        dunit = s.substr(sp + 1);
        size_t e = exp.find_first_of("eE");
        if (e != std::string::npos) {
            d.has_exp = true;
            d.pre = strtof(exp.substr(0, e).c_str(), nullptr);
            d.p = strtol(exp.substr(e + 1).c_str(), nullptr, 10);
        }
    } else {
        d.has_exp = true; // Some((1.0, 0))
        d.pre = 1.0f;
        d.p = 0;
    }
    // regex [/|*]?[A-Z|a-z]+\^?[0-9]*  (the '|' inside the classes is a literal)
    size_t i = 0;
    while (i < dunit.size()) {
        size_t st = i;
        char lead = 0;
        if (dunit[i] == '/' || dunit[i] == '|' || dunit[i] == '*') {
            if (i + 1 < dunit.size() && is_alpha_or_bar(dunit[i + 1])) {
                lead = dunit[i];
                ++i;
            } else if (dunit[i] == '|') {
                // '|' can itself start the letters class
            } else {
                ++i;
                continue;
            }
        }
        if (i >= dunit.size() || !is_alpha_or_bar(dunit[i])) {
            i = st + 1;
            continue;
        }
        size_t ks = i;
        while (i < dunit.size() && is_alpha_or_bar(dunit[i])) ++i;
        std::string key = dunit.substr(ks, i - ks);
        long p = (lead == '/') ? -1 : 1;
        bool has_caret = false;
        if (i < dunit.size() && dunit[i] == '^') {
            has_caret = true;
            ++i;
        }
        size_t ds = i;
        while (i < dunit.size() && dunit[i] >= '0' && dunit[i] <= '9') ++i;
        std::string digits = dunit.substr(ds, i - ds);
        if (has_caret) {
            p *= strtol(digits.c_str(), nullptr, 10); // the reference unwrap()s the parse; "" would panic there
        } else {
            key += digits; // without '^' the digits stay part of the key ("cm2")
        }
        if (lead == '|') key = std::string("|") + key; // s.chars().nth(0) is neither '/' nor '*': kept in the key
        d.insert(key, p);
    }
    return d;
}

static std::string derived_to_string(const Derived &d) {
    std::string ret;
    if (d.has_exp && !(d.pre == 1.0f && d.p == 0)) {
        if (d.pre - floorf(d.pre) == 0.0f) ret += fmt("%d", (int)d.pre);
        else ret += fmt("%g", d.pre);
        ret += fmt("E%ld ", d.p);
    }
    std::vector<std::pair<std::string, long>> v = d.derived;
    std::stable_sort(v.begin(), v.end(), [](const std::pair<std::string, long> &a, const std::pair<std::string, long> &b) { return a.second > b.second; });
    int count = 0;
    for (auto &kv : v) {
        if (kv.second == 0) continue;
        if (kv.second < 0) ret += "/";
        else if (kv.second > 0 && count > 0) ret += "*";
        ret += kv.first;
        long a = kv.second < 0 ? -kv.second : kv.second;
        if (a > 1) ret += fmt("^%ld", a);
        ++count;
    }
    return ret;
}

static std::string derived_combine(const std::string &a, const std::string &b, bool divide) {
    Derived l = derived_parse(a), r = derived_parse(b), o;
    if (l.has_exp && r.has_exp) {
        o.has_exp = true;
        o.pre = divide ? l.pre / r.pre : l.pre * r.pre;
        o.p = divide ? l.p - r.p : l.p + r.p;
    } else if (l.has_exp) {
        o.has_exp = true; o.pre = l.pre; o.p = l.p;
    } else if (r.has_exp) {
        o.has_exp = true; o.pre = r.pre; o.p = r.p;
    }
    o.derived = l.derived;
    for (auto &kv : r.derived) {
        long *e = o.find(kv.first);
        if (e) *e += divide ? -kv.second : kv.second;
        else o.derived.emplace_back(kv.first, divide ? -kv.second : kv.second);
    }
    return derived_to_string(o);
}

std::string unit_mul(bool has_a, const std::string &a, bool has_b, const std::string &b, bool *has_out) {
    if (has_a && has_b) { *has_out = true; return derived_combine(a, b, false); }
    *has_out = false; // (_, _) => Unit::None
    return "";
}
std::string unit_div(bool has_a, const std::string &a, bool has_b, const std::string &b, bool *has_out) {
    if (has_a && has_b) { *has_out = true; return derived_combine(a, b, true); }
    *has_out = false;
    return "";
}

// =============================================================================================================
// Precondition checks with the reference's wording (precond.rs)
// =============================================================================================================
Error unexpected(const std::string &m) { return make_error(AFK_ERR_UNEXPECTED_INPUT, m); }

Error try_into_unsigned(const char *name, int64_t v) { // precond.rs:3-15
    if (v >= 0) return ok();
    return unexpected(fmt("'%s' must be positive, but got %lld", name, (long long)v));
}

static Error has_gt_0_dim(const char *name, const afk_image *img, int64_t *frames) { // precond.rs:44-57
    if (img->ndim < 1) return unexpected(fmt("'%s' is a 0-dimensional image, cannot slice", name));
    *frames = img->shape[0];
    return ok();
}

static Error is_sliceable_range(const char *name, const afk_image *img, int64_t start, int64_t end) { // precond.rs:87-98
    int64_t frames = 0;
    if (Error e = has_gt_0_dim(name, img, &frames)) return e;
    if (!(start < end)) return unexpected(fmt("Expected start < end, but got %lld < %lld", (long long)start, (long long)end));
    if (!(end < frames + 1))
        return unexpected(fmt("'end' greater or equal to the input image '%s''s frame count (expected %lld < %lld)", name,
                              (long long)end, (long long)(frames + 1)));
    return ok();
}

static Error is_sliceable_frame(const char *name, const char *idx_name, const afk_image *img, int64_t frame) { // precond.rs:78-86
    int64_t frames = 0;
    if (Error e = has_gt_0_dim(name, img, &frames)) return e;
    if (!(frame < frames))
        return unexpected(fmt("'%s' greater or equal to the input image '%s''s frame count (expected %lld < %lld)", idx_name,
                              name, (long long)frame, (long long)frames));
    return ok();
}

static std::string dim_debug(const afk_image *img) { // {:?} of IxDyn: "[4563, 74, 74]"
    std::string s = "[";
    for (int i = 0; i < img->ndim; ++i) s += fmt(i ? ", %lld" : "%lld", (long long)img->shape[i]);
    return s + "]";
}

Error are_same_dim(const char *n1, const afk_image *a, const char *n2, const afk_image *b) { // precond.rs:101-114
    bool same = a->ndim == b->ndim;
    for (int i = 0; same && i < a->ndim; ++i) same = a->shape[i] == b->shape[i];
    if (same) return ok();
    return unexpected(fmt("Cannot compose images of different dimensions ('%s' has dimension %s, while '%s' has dimension %s)", n1,
                          dim_debug(a).c_str(), n2, dim_debug(b).c_str()));
}

Error same_device(const afk_image *a, int device) {
    if (a->device != device) return make_error(AFK_ERR_CUDA, fmt("CUDA: image lives on device %d but the calling thread uses device %d", a->device, device));
    return ok();
}

// WcsArray::make_slice with indices (0..ndim) (unit.rs:424-451): keep the first `ndim` axes, unit kept, tag dropped
static void meta_make_slice(const afk_image *src, afk_image *dst) {
    dst->meta.has_unit = src->meta.has_unit;
    dst->meta.unit = src->meta.unit;
    dst->meta.homogeneous = src->meta.homogeneous;
    dst->meta.has_tag = false;
    dst->meta.has_meta = src->meta.has_meta;
    if (src->meta.has_meta) {
        for (int i = 0; i < 4; ++i) dst->meta.axes[i] = (i < dst->ndim) ? src->meta.axes[i] : AxisMeta();
        // wcs.slice(&[0..ndim]) then transform(i, 0.0, 1.0): the kept FITS axes are unchanged, the dropped ones go
        // back to the identity {CRVAL 0, CRPIX 1, CDELT 1}
        dst->meta.wcs = src->meta.wcs;
        if (dst->meta.wcs.size() == 12)
            for (int i = dst->ndim; i < 4; ++i) {
                dst->meta.wcs[3 * i + 0] = 0.0;
                dst->meta.wcs[3 * i + 1] = 1.0;
                dst->meta.wcs[3 * i + 2] = 1.0;
            }
    }
}

// =============================================================================================================
// Operators on handles
// =============================================================================================================
Error op_reduce(const afk_image *cube, int nwin, const int64_t *starts, const int64_t *ends, int epilogue, const float *params,
                float ew_sign, bool want_minmax, afk_image **out, float *minmax) {
    cudaStream_t s;
    int dev;
    if (Error e = current_stream(&s, &dev)) return e;
    if (Error e = same_device(cube, dev)) return e;
    afk_image *o = nullptr;
    if (Error e = image_new(cube->ndim - 1, cube->shape + 1, &o)) return e;
    ReduceArgs a{};
    a.cube = cube->ptr;
    a.n2 = cube->frame_len();
    a.nwin = nwin;
    for (int w = 0; w < nwin; ++w) {
        a.start[w] = starts[w];
        a.end[w] = ends[w];
    }
    a.epilogue = epilogue;
    a.mode = effective_sum_mode();
    for (int i = 0; i < 4; ++i) a.p[i] = params ? params[i] : 0.0f;
    a.ew_sign = ew_sign;
    a.out = o->ptr;
    a.range_out = o->range; // range memo: a following image_min_max / logscale / display call needs no pass over the image
    Error err = ok();
    unsigned *mm = nullptr;
    do {
        if (a.n2 <= 0) break; // empty frames: nothing to launch
        if ((err = image_wait(cube, s))) break;
        if (want_minmax && (err = scratch_u32(s, 4, &mm))) break;
        a.minmax = mm;
        ReducePlan plan = reduce_plan(a, sm_count(dev));
        ReduceWorkspace ws;
        if ((err = reduce_workspace(plan, dev, s, &ws))) break;
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
        if (ce != cudaSuccess) {
            err = cuda_error(ce, "reduce_kernel launch");
            break;
        }
        image_note_read(cube, s);
        o->range_valid = o->range != nullptr;
        if (want_minmax) {
            unsigned h[4];
            ce = cudaMemcpyAsync(h, mm, sizeof h, cudaMemcpyDeviceToHost, s);
            if (ce == cudaSuccess) ce = cudaStreamSynchronize(s);
            if (ce != cudaSuccess) {
                err = cuda_error(ce, "minmax read-back");
                break;
            }
            // image_min_max (lib.rs:186-195): running min starts at f32::MAX, max at f32::MIN, NaN ignored
            float mn = h[2] ? minmax_decode_min(h[0]) : FLT_MAX, mx = h[2] ? minmax_decode_max(h[1]) : -FLT_MAX;
            minmax[0] = fminf(FLT_MAX, mn);
            minmax[1] = fmaxf(-FLT_MAX, mx);
        }
    } while (0);
    if (mm) scratch_free(s, mm);
    if (!err) err = image_mark_ready(o, s);
    if (err) {
        afk_image_release(o);
        return err;
    }
    *out = o;
    return ok();
}

Error op_map(int op, const afk_image *u, const afk_image *v, float a, float b, float c, int flag0, int flag1, afk_image **out) {
    cudaStream_t s;
    int dev;
    if (Error e = current_stream(&s, &dev)) return e;
    if (Error e = same_device(u, dev)) return e;
    if (v)
        if (Error e = same_device(v, dev)) return e;
    afk_image *o = nullptr;
    if (Error e = image_new(u->ndim, u->shape, &o)) return e;
    Error err = ok();
    do {
        if ((err = image_wait(u, s))) break;
        if (v && (err = image_wait(v, s))) break;
        MapArgs m{op, u->ptr, v ? v->ptr : nullptr, a, b, c, flag0, flag1, u->len(), o->ptr};
        cudaError_t ce = map_launch(m, sm_count(dev), s);
        if (ce != cudaSuccess) {
            err = cuda_error(ce, "map_kernel launch");
            break;
        }
        image_note_read(u, s);
        if (v) image_note_read(v, s);
        err = image_mark_ready(o, s);
    } while (0);
    if (err) {
        afk_image_release(o);
        return err;
    }
    *out = o;
    return ok();
}

Error op_min_max(const afk_image *img, float *mn, float *mx, bool *any) {
    cudaStream_t s;
    int dev;
    if (Error e = current_stream(&s, &dev)) return e;
    if (Error e = same_device(img, dev)) return e;
    // The three words {~key(min), key(max), any} reach the host through this thread's mapped read-back page: the kernel's last
    // CTA stores them there itself (or one asynchronous copy does, when the image carries a range memo), so the call is one
    // launch and one stream synchronisation -- no allocation, memset or copy around the kernel.
    char *page = nullptr;
    if (Error e = readback_page(&page)) return e;
    volatile unsigned *h = reinterpret_cast<volatile unsigned *>(page + 512);
    const bool memo = img->range_valid && img->range; // left by the kernel that wrote the pixels: no pass over the image
    h[0] = h[1] = h[2] = 0u; // an empty image launches nothing
    if (Error e = image_wait(img, s)) return e;
    cudaError_t ce = cudaSuccess;
    StreamScratch &sc = tls().scratch[std::make_pair(dev, s)];
    if (memo) {
        ce = cudaMemcpyAsync(page + 512, img->range, 3 * sizeof(unsigned), cudaMemcpyDeviceToHost, s);
    } else if (img->len() > 0) {
        if (!sc.minmax_self) {
            ce = cudaMallocAsync(reinterpret_cast<void **>(&sc.minmax_self), 4 * sizeof(unsigned), s);
            if (ce == cudaSuccess) ce = cudaMemsetAsync(sc.minmax_self, 0, 4 * sizeof(unsigned), s);
            else sc.minmax_self = nullptr;
        }
        if (ce == cudaSuccess) ce = minmax_launch_readback(img->ptr, img->len(), sc.minmax_self, reinterpret_cast<unsigned *>(page + 512), sm_count(dev), s);
    }
    if (ce == cudaSuccess) ce = cudaStreamSynchronize(s);
    if (ce != cudaSuccess) {
        // the words may not have been returned to zero: start from a fresh block next time
        if (sc.minmax_self) cudaFreeAsync(sc.minmax_self, s);
        sc.minmax_self = nullptr;
        return cuda_error(ce, "minmax_kernel");
    }
    image_note_read(img, s);
    const unsigned h0 = h[0], h1 = h[1], h2 = h[2];
    *any = h2 != 0;
    *mn = *any ? minmax_decode_min(h0) : std::numeric_limits<float>::quiet_NaN();
    *mx = *any ? minmax_decode_max(h1) : std::numeric_limits<float>::quiet_NaN();
    return ok();
}

// clone() of a WcsArray keeps meta, unit and tag
static void meta_clone(const afk_image *src, afk_image *dst) { dst->meta = src->meta; }

// ---- result helpers ---------------------------------------------------------------------------------------------
void set_ok(afk_result *r, afk_value *v) {
    r->status = AFK_OK;
    r->value = v;
    r->message = nullptr;
}
void set_err(afk_result *r, const Error &e) {
    r->status = e.status;
    r->value = nullptr;
    r->message = strdup(e.message.c_str());
    set_last_error(e.message);
}
void set_image(afk_result *r, afk_image *img) {
    afk_value *v = afk_value_image(img);
    afk_image_release(img); // the value holds the only reference
    set_ok(r, v);
}

// cake_fn! wrapper (aflak_cake/src/lib.rs:129-141): mismatching variants are a panic in the reference
bool args_match(const afk_value *const *in, size_t n_in, std::initializer_list<afk_type> types, afk_result *out, size_t n_out) {
    bool good = n_in == types.size();
    size_t i = 0;
    for (afk_type t : types) {
        if (!good) break;
        good = in[i] && in[i]->type == t;
        ++i;
    }
    if (!good)
        for (size_t k = 0; k < n_out; ++k) set_err(&out[k], make_error(AFK_ERR_PANIC, "Unexpected argument!"));
    return good;
}

#define AFK_OP(name) static afk_status name(const afk_value *const *in, size_t n_in, afk_result *out, size_t n_out)
#define AFK_NEED_OUT(n)                                                                            \
    if (n_out != (n)) return fail(make_error(AFK_ERR_PANIC, "wrong number of output slots"));

// ---- integral / average (lib.rs:384-411, 1303-1330) --------------------------------------------------------------
static afk_status reduce_op(int epilogue, const afk_value *const *in, size_t n_in, afk_result *out, size_t n_out) {
    AFK_NEED_OUT(3)
    if (!args_match(in, n_in, {AFK_T_IMAGE, AFK_T_INTEGER, AFK_T_INTEGER}, out, n_out)) return AFK_ERR_PANIC;
    const afk_image *img = in[0]->img;
    const int64_t start = in[1]->i, end = in[2]->i;
    // second and third outputs (lib.rs:393-395): pure host arithmetic in f32
    const float middle = ((float)start + (float)end) / 2.0f;
    const float width = (float)end - (float)start;
    Error e = try_into_unsigned("start", start);
    if (!e) e = try_into_unsigned("end", end);
    if (!e) e = is_sliceable_range("image", img, start, end);
    afk_image *res = nullptr;
    if (!e) e = op_reduce(img, 1, &start, &end, epilogue, nullptr, 1.0f, false, &res, nullptr);
    if (e) set_err(&out[0], e);
    else {
        meta_make_slice(img, res);
        set_image(&out[0], res);
    }
    set_ok(&out[1], afk_value_float(middle));
    set_ok(&out[2], afk_value_float(width));
    return out[0].status;
}
AFK_OP(op_integral) { return reduce_op(EPI_SUM, in, n_in, out, n_out); }
AFK_OP(op_average) { return reduce_op(EPI_MEAN, in, n_in, out, n_out); }

// ---- slice_one_frame (lib.rs:352-359, 912-925): zero-copy view that keeps the cube alive --------------------------
AFK_OP(op_slice_one_frame) {
    AFK_NEED_OUT(1)
    if (!args_match(in, n_in, {AFK_T_IMAGE, AFK_T_INTEGER}, out, n_out)) return AFK_ERR_PANIC;
    afk_image *img = in[0]->img;
    const int64_t frame = in[1]->i;
    Error e = try_into_unsigned("frame_idx", frame);
    if (!e) e = is_sliceable_frame("input_img", "frame_idx", img, frame);
    if (e) {
        set_err(&out[0], e);
        return e.status;
    }
    afk_image *view = new afk_image();
    view->ndim = img->ndim - 1;
    for (int i = 0; i < view->ndim; ++i) view->shape[i] = img->shape[i + 1];
    view->ptr = img->ptr + frame * img->frame_len();
    view->bytes = (size_t)view->len() * sizeof(float);
    view->owned = false;
    view->parent = img;
    afk_image_retain(img);
    view->device = img->device;
    view->stream = img->stream;
    if (img->ready) { // the view is ready when its parent is
        view->ready = event_get();
        cudaStream_t s;
        if (!current_stream(&s, nullptr)) {
            cudaStreamWaitEvent(s, img->ready, 0);
            cudaEventRecord(view->ready, s);
            view->stream = s;
        }
    }
    meta_make_slice(img, view);
    set_image(&out[0], view);
    return AFK_OK;
}

// ---- range_specification (lib.rs:171-179, 1229-1231): sub-cube copy -----------------------------------------------
AFK_OP(op_range_specification) {
    AFK_NEED_OUT(1)
    if (!args_match(in, n_in, {AFK_T_IMAGE, AFK_T_INTEGER, AFK_T_INTEGER}, out, n_out)) return AFK_ERR_PANIC;
    const afk_image *img = in[0]->img;
    const int64_t start = in[1]->i, end = in[2]->i;
    Error e = try_into_unsigned("start", start);
    if (!e) e = try_into_unsigned("end", end);
    if (!e) e = is_sliceable_range("image", img, start, end);
    afk_image *res = nullptr;
    cudaStream_t s = nullptr;
    if (!e) e = current_stream(&s, nullptr);
    if (!e) {
        int64_t shape[AFK_MAX_NDIM];
        for (int i = 0; i < img->ndim; ++i) shape[i] = img->shape[i];
        shape[0] = end - start;
        e = image_new(img->ndim, shape, &res);
    }
    if (!e) e = image_wait(img, s);
    if (!e && res->bytes) {
        cudaError_t ce = cudaMemcpyAsync(res->ptr, img->ptr + start * img->frame_len(), res->bytes, cudaMemcpyDeviceToDevice, s);
        if (ce != cudaSuccess) e = cuda_error(ce, "cudaMemcpyAsync(D2D)");
    }
    if (!e) {
        image_note_read(img, s);
        e = image_mark_ready(res, s);
    }
    if (e) {
        if (res) afk_image_release(res);
        set_err(&out[0], e);
        return e.status;
    }
    meta_make_slice(img, res);
    set_image(&out[0], res);
    return AFK_OK;
}

// ---- image_min_max (lib.rs:180-197) ----------------------------------------------------------------------------------
AFK_OP(op_image_min_max) {
    AFK_NEED_OUT(2)
    if (!args_match(in, n_in, {AFK_T_IMAGE}, out, n_out)) return AFK_ERR_PANIC;
    float mn, mx;
    bool any;
    if (Error e = op_min_max(in[0]->img, &mn, &mx, &any)) {
        set_err(&out[0], e);
        set_err(&out[1], e);
        return e.status;
    }
    // min starts at f32::MAX, max at f32::MIN; Rust's min/max skip a NaN operand
    set_ok(&out[0], afk_value_float(any ? fminf(FLT_MAX, mn) : FLT_MAX));
    set_ok(&out[1], afk_value_float(any ? fmaxf(-FLT_MAX, mx) : -FLT_MAX));
    return AFK_OK;
}

// ---- ratio_from_bands (lib.rs:226-246): host scalars ------------------------------------------------------------------
std::string rust_f32_display(float f) {
    // Rust's `{}` for f32 prints the shortest string that round-trips, without exponent for these magnitudes
    if (f != f) return "NaN";
    if (std::isinf(f)) return f > 0 ? "inf" : "-inf";
    if (f == floorf(f) && fabsf(f) < 1e16f) return fmt("%.0f", (double)f);
    for (int prec = 1; prec < 12; ++prec) {
        std::string s = fmt("%.*g", prec, (double)f);
        if (strtof(s.c_str(), nullptr) == f && s.find('e') == std::string::npos) return s;
    }
    return fmt("%.9g", (double)f);
}

AFK_OP(op_ratio_from_bands) {
    AFK_NEED_OUT(2)
    if (!args_match(in, n_in, {AFK_T_FLOAT, AFK_T_FLOAT, AFK_T_FLOAT}, out, n_out)) return AFK_ERR_PANIC;
    const float z = in[0]->f[0], z1 = in[1]->f[0], z2 = in[2]->f[0];
    if (!(z1 < z && z < z2)) {
        Error e = unexpected(fmt("wrong magnitude correlation (%s < %s < %s)", rust_f32_display(z1).c_str(),
                                 rust_f32_display(z).c_str(), rust_f32_display(z2).c_str()));
        set_err(&out[0], e);
        set_err(&out[1], e);
        return e.status;
    }
    const float off_ratio_1 = (z - z1) / (z2 - z1);
    const float off_ratio_2 = 1.0f - off_ratio_1;
    set_ok(&out[0], afk_value_float(off_ratio_2));
    set_ok(&out[1], afk_value_float(off_ratio_1));
    return AFK_OK;
}

// ---- unary elementwise operators that clone the input (meta, unit and tag kept) ---------------------------------------
static afk_status unary_clone_op(int op, const afk_image *img, float a, float b, float c, int f0, int f1, afk_result *out) {
    afk_image *res = nullptr;
    if (Error e = op_map(op, img, nullptr, a, b, c, f0, f1, &res)) {
        set_err(&out[0], e);
        return e.status;
    }
    meta_clone(img, res);
    set_image(&out[0], res);
    return AFK_OK;
}

AFK_OP(op_convert_to_logscale) { // lib.rs:282-290, 1598-1611
    AFK_NEED_OUT(1)
    if (!args_match(in, n_in, {AFK_T_IMAGE, AFK_T_FLOAT, AFK_T_FLOAT, AFK_T_FLOAT}, out, n_out)) return AFK_ERR_PANIC;
    return unary_clone_op(MAP_LOGSCALE, in[0]->img, in[1]->f[0], in[2]->f[0], in[3]->f[0], 0, 0, out);
}
AFK_OP(op_log10) { // lib.rs:291-300, 1613-1617
    AFK_NEED_OUT(1)
    if (!args_match(in, n_in, {AFK_T_IMAGE}, out, n_out)) return AFK_ERR_PANIC;
    return unary_clone_op(MAP_LOG10, in[0]->img, 0, 0, 0, 0, 0, out);
}
AFK_OP(op_negation) { // lib.rs:301-308, 1640-1645
    AFK_NEED_OUT(1)
    if (!args_match(in, n_in, {AFK_T_IMAGE}, out, n_out)) return AFK_ERR_PANIC;
    return unary_clone_op(MAP_NEGATION, in[0]->img, 0, 0, 0, 0, 0, out);
}
AFK_OP(op_normalize_image) { // lib.rs:266-274
    AFK_NEED_OUT(1)
    if (!args_match(in, n_in, {AFK_T_IMAGE, AFK_T_FLOAT, AFK_T_FLOAT}, out, n_out)) return AFK_ERR_PANIC;
    return unary_clone_op(MAP_NORMALIZE, in[0]->img, in[1]->f[0], in[2]->f[0], 0, 0, 0, out);
}
AFK_OP(op_clip_image) { // lib.rs:257-265, 1258-1274
    AFK_NEED_OUT(1)
    if (!args_match(in, n_in, {AFK_T_IMAGE, AFK_T_FLOAT, AFK_T_BOOL, AFK_T_FLOAT, AFK_T_BOOL}, out, n_out)) return AFK_ERR_PANIC;
    return unary_clone_op(MAP_CLIP, in[0]->img, in[1]->f[0], in[3]->f[0], 0, in[2]->b, in[4]->b, out);
}
AFK_OP(op_replace_nan_image) { // lib.rs:275-281, 1276-1286
    AFK_NEED_OUT(1)
    if (!args_match(in, n_in, {AFK_T_IMAGE, AFK_T_FLOAT}, out, n_out)) return AFK_ERR_PANIC;
    return unary_clone_op(MAP_REPLACE_NAN, in[0]->img, in[1]->f[0], 0, 0, 0, 0, out);
}

AFK_OP(op_change_tag) { // lib.rs:325-333, 1788-1793: metadata only, the device buffer is shared
    AFK_NEED_OUT(1)
    if (!args_match(in, n_in, {AFK_T_IMAGE, AFK_T_STR}, out, n_out)) return AFK_ERR_PANIC;
    afk_image *img = in[0]->img;
    afk_image *view = new afk_image();
    view->ndim = img->ndim;
    for (int i = 0; i < img->ndim; ++i) view->shape[i] = img->shape[i];
    view->ptr = img->ptr;
    view->bytes = img->bytes;
    view->owned = false;
    view->parent = img;
    afk_image_retain(img);
    view->device = img->device;
    view->stream = img->stream;
    if (img->ready) {
        cudaStream_t s;
        if (!current_stream(&s, nullptr)) {
            view->ready = event_get();
            cudaStreamWaitEvent(s, img->ready, 0);
            cudaEventRecord(view->ready, s);
            view->stream = s;
        }
    }
    view->meta = img->meta;
    view->meta.has_tag = true;
    view->meta.tag = in[1]->s;
    set_image(&out[0], view);
    return AFK_OK;
}

// ---- apply_tone_curve (lib.rs:309-316, 1858-1871) --------------------------------------------------------------------
AFK_OP(op_apply_tone_curve) {
    AFK_NEED_OUT(1)
    if (!args_match(in, n_in, {AFK_T_IMAGE, AFK_T_TONECURVE}, out, n_out)) return AFK_ERR_PANIC;
    const afk_image *img = in[0]->img;
    const std::vector<float> &table = in[1]->arr;
    Error e = ok();
    if (table.empty()) e = make_error(AFK_ERR_PANIC, "attempt to subtract with overflow (empty tone-curve table)");
    else if (table.size() > (size_t)kMaxToneTable) e = unexpected(fmt("tone-curve table too long (%zu > %d)", table.size(), kMaxToneTable));
    cudaStream_t s = nullptr;
    int dev = 0;
    if (!e) e = current_stream(&s, &dev);
    if (!e) e = same_device(img, dev);
    afk_image *res = nullptr;
    float *dtab = nullptr;
    unsigned *flag = nullptr;
    do {
        if (e) break;
        if ((e = image_new(img->ndim, img->shape, &res))) break;
        if ((e = image_wait(img, s))) break;
        if ((e = panic_word_get(dev, &flag))) break;
        tls().panic_tone_len = table.size();
        cudaError_t ce = cudaMallocAsync(reinterpret_cast<void **>(&dtab), table.size() * sizeof(float), s);
        if (ce == cudaSuccess) ce = cudaMemcpyAsync(dtab, table.data(), table.size() * sizeof(float), cudaMemcpyHostToDevice, s);
        if (ce == cudaSuccess) ce = tone_curve_launch(img->ptr, img->len(), dtab, (int)table.size(), res->ptr, flag, sm_count(dev), s);
        // the operator returns Ok or the reference's panic: the index check needs the kernel to have finished (one stream
        // synchronisation; the flag itself is a word of mapped host memory, no copy)
        if (ce == cudaSuccess) ce = cudaStreamSynchronize(s);
        if (ce != cudaSuccess) {
            e = cuda_error(ce, "tone_kernel");
            break;
        }
        image_note_read(img, s);
        if ((e = panic_word_take(dev))) break;
        e = image_mark_ready(res, s);
    } while (0);
    if (dtab) cudaFreeAsync(dtab, s);
    if (e) {
        if (res) afk_image_release(res);
        set_err(&out[0], e);
        return e.status;
    }
    // WcsArray::from_array(Dimensioned::new(arr, Unit::None)): meta, unit and tag dropped
    set_image(&out[0], res);
    return AFK_OK;
}

// ---- linear_composition (lib.rs:334-342, 1288-1297) --------------------------------------------------------------------
AFK_OP(op_linear_composition) {
    AFK_NEED_OUT(1)
    if (!args_match(in, n_in, {AFK_T_IMAGE, AFK_T_IMAGE, AFK_T_FLOAT, AFK_T_FLOAT}, out, n_out)) return AFK_ERR_PANIC;
    const afk_image *u = in[0]->img, *v = in[1]->img;
    Error e = are_same_dim("i1", u, "i2", v);
    afk_image *res = nullptr;
    if (!e) e = op_map(MAP_LINCOMP, u, v, in[2]->f[0], in[3]->f[0], 0, 0, 0, &res);
    if (e) {
        set_err(&out[0], e);
        return e.status;
    }
    // (&i1 * a) keeps i1's meta/unit; WcsArray + WcsArray (unit.rs:817-832): meta if equal else None, unit of lhs,
    // homogeneous only if the units agree (unit.rs:659-673); tag dropped
    res->meta.has_meta = u->meta.meta_equal(v->meta) && u->meta.has_meta;
    if (res->meta.has_meta) {
        for (int i = 0; i < 4; ++i) res->meta.axes[i] = u->meta.axes[i];
        res->meta.wcs = u->meta.wcs;
    }
    res->meta.has_unit = u->meta.has_unit;
    res->meta.unit = u->meta.unit;
    const bool same_unit = u->meta.has_unit == v->meta.has_unit && (!u->meta.has_unit || u->meta.unit == v->meta.unit);
    res->meta.homogeneous = same_unit && u->meta.homogeneous && v->meta.homogeneous;
    set_image(&out[0], res);
    return AFK_OK;
}

// ---- ratio: &WcsArray / &WcsArray (unit.rs:863-880); additive registry entry, see SURVEY.md 2a "ratio note" ---------------
static afk_status ratio_impl(const afk_image *u, const afk_image *v, afk_result *out) {
    Error e = are_same_dim("i1", u, "i2", v);
    afk_image *res = nullptr;
    if (!e) e = op_map(MAP_RATIO, u, v, 0, 0, 0, 0, 0, &res);
    if (e) {
        set_err(&out[0], e);
        return e.status;
    }
    res->meta.has_meta = u->meta.has_meta;
    for (int i = 0; i < 4; ++i) res->meta.axes[i] = u->meta.axes[i];
    res->meta.wcs = u->meta.wcs;
    res->meta.unit = unit_div(u->meta.has_unit, u->meta.unit, v->meta.has_unit, v->meta.unit, &res->meta.has_unit);
    res->meta.homogeneous = true;
    set_image(&out[0], res);
    return AFK_OK;
}
AFK_OP(op_ratio) {
    AFK_NEED_OUT(1)
    if (!args_match(in, n_in, {AFK_T_IMAGE, AFK_T_IMAGE}, out, n_out)) return AFK_ERR_PANIC;
    return ratio_impl(in[0]->img, in[1]->img, out);
}

// ---- image_multiplier (lib.rs:343-351, 1619-1638): powf(u,a) * powf(v,b) ---------------------------------------------------
AFK_OP(op_image_multiplier) {
    AFK_NEED_OUT(1)
    if (!args_match(in, n_in, {AFK_T_IMAGE, AFK_T_IMAGE, AFK_T_FLOAT, AFK_T_FLOAT}, out, n_out)) return AFK_ERR_PANIC;
    const afk_image *u = in[0]->img, *v = in[1]->img;
    Error e = are_same_dim("i1", u, "i2", v);
    afk_image *res = nullptr;
    if (!e) e = op_map(MAP_MULTIPLIER, u, v, in[2]->f[0], in[3]->f[0], 0, 0, 0, &res);
    if (e) {
        set_err(&out[0], e);
        return e.status;
    }
    // WcsArray * WcsArray (unit.rs:901-918): meta of lhs, units multiplied, tag dropped
    res->meta.has_meta = u->meta.has_meta;
    for (int i = 0; i < 4; ++i) res->meta.axes[i] = u->meta.axes[i];
    res->meta.wcs = u->meta.wcs;
    res->meta.unit = unit_mul(u->meta.has_unit, u->meta.unit, v->meta.has_unit, v->meta.unit, &res->meta.has_unit);
    res->meta.homogeneous = true;
    set_image(&out[0], res);
    return AFK_OK;
}

// ---- create_equivalent_width (lib.rs:498-508, 1584-1596) ---------------------------------------------------------------------
static void ew_unit(const afk_image *off, afk_image *res) {
    // (off - on) keeps off's unit; * Dimensioned(fl, "Ang") multiplies units; / off divides by off's unit
    bool h1, h2;
    std::string u1 = unit_mul(off->meta.has_unit, off->meta.unit, true, "Ang", &h1);
    std::string u2 = unit_div(h1, u1, off->meta.has_unit, off->meta.unit, &h2);
    res->meta = ImageMeta();
    res->meta.has_unit = h2;
    res->meta.unit = u2;
    res->meta.homogeneous = true;
}

AFK_OP(op_create_equivalent_width) {
    AFK_NEED_OUT(1)
    if (!args_match(in, n_in, {AFK_T_IMAGE, AFK_T_IMAGE, AFK_T_FLOAT, AFK_T_FLOAT, AFK_T_BOOL}, out, n_out)) return AFK_ERR_PANIC;
    const afk_image *off = in[0]->img, *on = in[1]->img;
    // the reference relies on ndarray broadcasting rules and panics on incompatible shapes; equal shapes are required here
    Error e = are_same_dim("i_off", off, "i_on", on);
    afk_image *res = nullptr;
    if (!e) e = op_map(MAP_EW, off, on, in[2]->f[0], in[3]->f[0], in[4]->b ? -1.0f : 1.0f, 0, 0, &res);
    if (e) {
        set_err(&out[0], e);
        return e.status;
    }
    ew_unit(off, res);
    set_image(&out[0], res);
    return AFK_OK;
}

AFK_OP(op_create_emission_line_map) { // lib.rs:527-537, 1770-1786
    AFK_NEED_OUT(1)
    if (!args_match(in, n_in, {AFK_T_IMAGE, AFK_T_IMAGE, AFK_T_FLOAT, AFK_T_FLOAT, AFK_T_BOOL}, out, n_out)) return AFK_ERR_PANIC;
    const afk_image *off = in[0]->img, *on = in[1]->img;
    Error e = are_same_dim("i_off", off, "i_on", on);
    afk_image *res = nullptr;
    if (!e) e = op_map(MAP_EMISSION_LINE, off, on, in[2]->f[0], in[3]->f[0], in[4]->b ? -1.0f : 1.0f, 0, 0, &res);
    if (e) {
        set_err(&out[0], e);
        return e.status;
    }
    res->meta = ImageMeta(); // from_array(.., Unit::None)
    set_image(&out[0], res);
    return AFK_OK;
}

// ---- down_sampling (lib.rs:554-561, 1894-1962) -----------------------------------------------------------------------------------
AFK_OP(op_down_sampling) {
    AFK_NEED_OUT(1)
    if (!args_match(in, n_in, {AFK_T_IMAGE, AFK_T_INTEGER}, out, n_out)) return AFK_ERR_PANIC;
    const afk_image *img = in[0]->img;
    const int64_t n = in[1]->i;
    Error e = ok();
    if (!(n > 0)) e = unexpected(fmt("n should be greater than 0, %lld > 0", (long long)n));
    else if (img->ndim != 2 && img->ndim != 3) e = unexpected("This is neither 2 nor 3 dimensional image!");
    cudaStream_t s = nullptr;
    int dev = 0;
    if (!e) e = current_stream(&s, &dev);
    if (!e) e = same_device(img, dev);
    afk_image *res = nullptr;
    if (!e) {
        const int64_t c = img->ndim == 3 ? img->shape[0] : 1;
        const int64_t ny = img->shape[img->ndim - 2], nx = img->shape[img->ndim - 1];
        const int64_t oy = ny / n + (ny % n != 0), ox = nx / n + (nx % n != 0);
        int64_t shape[3];
        if (img->ndim == 3) { shape[0] = c; shape[1] = oy; shape[2] = ox; }
        else { shape[0] = oy; shape[1] = ox; }
        e = image_new(img->ndim, shape, &res);
        if (!e) e = image_wait(img, s);
        if (!e) {
            cudaError_t ce = downsample_launch(img->ptr, c, ny, nx, n, res->ptr, oy, ox, s);
            if (ce != cudaSuccess) e = cuda_error(ce, "downsample_kernel launch");
        }
        if (!e) {
            image_note_read(img, s);
            e = image_mark_ready(res, s);
        }
    }
    if (e) {
        if (res) afk_image_release(res);
        set_err(&out[0], e);
        return e.status;
    }
    // from_array_and_tag(Dimensioned::new(out, Unit::None), tag): unit and WCS dropped, tag kept
    res->meta = ImageMeta();
    res->meta.has_tag = img->meta.has_tag;
    res->meta.tag = img->meta.tag;
    set_image(&out[0], res);
    return AFK_OK;
}

// =============================================================================================================
// Other axis-0 reductions: variance, stddev, median, arg maps, centrobaric wavelength
// =============================================================================================================

// linear WCS of one FITS axis: meta.wcs holds {CRVAL, CRPIX, CDELT} for each of the 4 axes (fitrs WCS restated as
// the per-axis linear transform; parity unpinned); an image whose header carried no WCS keys has an empty
// blob and maps a pixel to itself
static bool wcs_axis(const afk_image *img, int axis, double *crval, double *crpix, double *cdelt) {
    if (!img->meta.has_meta || img->meta.wcs.size() != 12 || axis < 0 || axis > 3) return false;
    *crval = img->meta.wcs[3 * axis + 0];
    *crpix = img->meta.wcs[3 * axis + 1];
    *cdelt = img->meta.wcs[3 * axis + 2];
    return true;
}

float image_pix2world(const afk_image *img, int axis, float pixel) { // unit.rs:335-342
    double crval, crpix, cdelt;
    if (!wcs_axis(img, axis, &crval, &crpix, &cdelt)) return pixel;
    return (float)(crval + cdelt * ((double)pixel + (1.0 - crpix)));
}

// the axis run_create_argmap / run_centroid convert indices with (lib.rs:1413-1421): the last axis named "WAVE" whose
// unit is Angstrom / Ang, else axis 0 (and then no unit on the result)
static int wave_axis(const afk_image *img, bool *found) {
    int trypix = 0;
    *found = false;
    for (int n = 0; n < 4; ++n) {
        const AxisMeta &a = img->meta.axes[n];
        if (a.name == "WAVE" && (a.unit == "Angstrom" || a.unit == "Ang")) {
            trypix = n;
            *found = true;
        }
    }
    return trypix;
}

static Error op_axis0(const afk_image *cube, Axis0Args a, const afk_image *m_lo, const afk_image *m_hi, afk_image **out,
                      afk_image **out2, const char *panic_message) {
    cudaStream_t s;
    int dev;
    if (Error e = current_stream(&s, &dev)) return e;
    if (Error e = same_device(cube, dev)) return e;
    afk_image *o = nullptr, *o2 = nullptr;
    if (Error e = image_new(cube->ndim - 1, cube->shape + 1, &o)) return e;
    if (out2) {
        if (Error e = image_new(cube->ndim - 1, cube->shape + 1, &o2)) {
            afk_image_release(o);
            return e;
        }
    }
    a.mode = effective_sum_mode();
    a.cube = cube->ptr;
    a.n2 = cube->frame_len();
    a.nz = cube->shape[0];
    a.mask_lo = m_lo ? m_lo->ptr : nullptr;
    a.mask_hi = m_hi ? m_hi->ptr : nullptr;
    a.out = o->ptr;
    a.out2 = o2 ? o2->ptr : nullptr;
    Error err = ok();
    void *partials = nullptr;
    unsigned *flag = nullptr;
    do {
        if (a.n2 <= 0) break;
        if ((err = image_wait(cube, s))) break;
        if (m_lo && (err = image_wait(m_lo, s))) break;
        if (m_hi && (err = image_wait(m_hi, s))) break;
        cudaError_t ce = cudaSuccess;
        if (a.op == A0_MEDIAN_OP) {
            if ((err = scratch_u32(s, kMedianScratchWords, &flag))) break;
            ce = median_launch(a.cube, a.n2, a.start, a.end, a.out, flag, s);
        } else {
            Axis0Plan plan = a.op == A0_CENTROID_MASK ? Axis0Plan{1, 256, 0, 1, 0, 0} : axis0_plan(a, sm_count(dev));
            if (plan.partial_bytes) {
                ce = cudaMallocAsync(&partials, plan.partial_bytes, s);
                if (ce != cudaSuccess) {
                    err = cuda_error(ce, "cudaMallocAsync(axis0 partials)");
                    break;
                }
            }
            if (a.op == A0_CENTROID_MASK && (err = scratch_u32(s, 1, &flag))) break;
            ce = axis0_launch(a, plan, partials, flag, s);
        }
        if (ce != cudaSuccess) {
            err = cuda_error(ce, "axis0 kernel launch");
            break;
        }
        image_note_read(cube, s);
        if (m_lo) image_note_read(m_lo, s);
        if (m_hi) image_note_read(m_hi, s);
        if (flag) { // the reference panics on these inputs: one 4-byte sync, like the tone-curve index check
            unsigned h = 0;
            ce = cudaMemcpyAsync(&h, flag, sizeof h, cudaMemcpyDeviceToHost, s);
            if (ce == cudaSuccess) ce = cudaStreamSynchronize(s);
            if (ce != cudaSuccess) {
                err = cuda_error(ce, "axis0 flag read-back");
                break;
            }
            if (h) {
                err = make_error(AFK_ERR_PANIC, panic_message);
                break;
            }
        }
    } while (0);
    if (partials) cudaFreeAsync(partials, s);
    if (flag) scratch_free(s, flag);
    if (!err) err = image_mark_ready(o, s);
    if (!err && o2) err = image_mark_ready(o2, s);
    if (err) {
        afk_image_release(o);
        if (o2) afk_image_release(o2);
        return err;
    }
    *out = o;
    if (out2) *out2 = o2;
    return ok();
}

static Error range_checks(const afk_image *img, int64_t start, int64_t end) {
    Error e = try_into_unsigned("start", start);
    if (!e) e = try_into_unsigned("end", end);
    if (!e) e = is_sliceable_range("image", img, start, end);
    return e;
}

// ---- variance / stddev (lib.rs:412-427, 1332-1338) ---------------------------------------------------------------
static afk_status var_op(int op, const afk_value *const *in, size_t n_in, afk_result *out, size_t n_out) {
    AFK_NEED_OUT(1)
    if (!args_match(in, n_in, {AFK_T_IMAGE, AFK_T_INTEGER, AFK_T_INTEGER}, out, n_out)) return AFK_ERR_PANIC;
    const afk_image *img = in[0]->img;
    const int64_t start = in[1]->i, end = in[2]->i;
    Error e = range_checks(img, start, end);
    // ndarray 0.12.1 var_axis panics when ddof (1.0) >= the number of frames
    if (!e && end - start < 2)
        e = make_error(AFK_ERR_PANIC, "`ddof` needs to be strictly smaller than the length of the axis you are computing the variance/standard deviation for!");
    afk_image *res = nullptr;
    if (!e) {
        Axis0Args a{};
        a.op = op;
        a.start = start;
        a.end = end;
        e = op_axis0(img, a, nullptr, nullptr, &res, nullptr, "");
    }
    if (e) {
        set_err(&out[0], e);
        return e.status;
    }
    meta_make_slice(img, res);
    set_image(&out[0], res);
    return AFK_OK;
}
AFK_OP(op_variance) { return var_op(A0_VARIANCE, in, n_in, out, n_out); }
AFK_OP(op_stddev) { return var_op(A0_STDDEV, in, n_in, out, n_out); }

// ---- median (lib.rs:428-435, 1340-1372) ------------------------------------------------------------------------------
AFK_OP(op_median) {
    AFK_NEED_OUT(1)
    if (!args_match(in, n_in, {AFK_T_IMAGE, AFK_T_INTEGER, AFK_T_INTEGER}, out, n_out)) return AFK_ERR_PANIC;
    const afk_image *img = in[0]->img;
    const int64_t start = in[1]->i, end = in[2]->i;
    Error e = range_checks(img, start, end);
    afk_image *res = nullptr;
    if (!e) {
        Axis0Args a{};
        a.op = A0_MEDIAN_OP;
        a.start = start;
        a.end = end;
        // vals.sort_by(|a, b| a.partial_cmp(b).unwrap()) panics on the first NaN comparison
        e = op_axis0(img, a, nullptr, nullptr, &res, nullptr, "called `Option::unwrap()` on a `None` value (NaN in median)");
    }
    if (e) {
        set_err(&out[0], e);
        return e.status;
    }
    // WcsArray::from_array(Dimensioned::new(result, Unit::None)): no meta, no unit, no tag
    set_image(&out[0], res);
    return AFK_OK;
}

// ---- arg maps (lib.rs:436-458, 1374-1464) -----------------------------------------------------------------------------
// meta of the *cube* is attached unchanged to the 2-D result ("FIXME: Handle metadata" in the reference), tag kept
static void argmap_meta(const afk_image *img, afk_image *res, bool wave_found) {
    res->meta = img->meta;
    res->meta.has_unit = wave_found;
    res->meta.unit = wave_found ? "Ang" : "";
    res->meta.homogeneous = true;
}

static afk_status argmap_op(const afk_image *img, int64_t start, int64_t end, bool is_min, bool actual, int64_t r1, int64_t r2,
                            int out2_kind, afk_result *out) {
    Error e = range_checks(img, start, end);
    // image.meta().clone().unwrap() (lib.rs:1413)
    if (!e && !img->meta.has_meta) e = make_error(AFK_ERR_PANIC, "called `Option::unwrap()` on a `None` value");
    afk_image *res = nullptr, *res2 = nullptr;
    bool found = false;
    if (!e) {
        const int trypix = wave_axis(img, &found);
        Axis0Args a{};
        a.op = A0_ARGFOLD;
        a.start = start;
        a.end = end;
        a.is_min = is_min;
        a.r1 = r1;
        a.r2 = r2;
        a.out2_kind = out2_kind;
        a.use_wcs = actual && wcs_axis(img, trypix, &a.crval, &a.crpix, &a.cdelt);
        e = op_axis0(img, a, nullptr, nullptr, &res, &res2, "");
    }
    if (e) {
        set_err(&out[0], e);
        set_err(&out[1], e);
        return e.status;
    }
    argmap_meta(img, res, found);
    if (out2_kind == 2) argmap_meta(img, res2, found);
    else meta_make_slice(img, res2); // run_minmax goes through reduce_array_slice
    set_image(&out[0], res);
    set_image(&out[1], res2);
    return AFK_OK;
}

AFK_OP(op_extract_argmin_max_wavelength) {
    AFK_NEED_OUT(2)
    if (!args_match(in, n_in, {AFK_T_IMAGE, AFK_T_INTEGER, AFK_T_INTEGER, AFK_T_BOOL}, out, n_out)) return AFK_ERR_PANIC;
    return argmap_op(in[0]->img, in[1]->i, in[2]->i, in[3]->b, true, 0, 0, 1, out);
}
AFK_OP(op_peak_based_wavelength_range) {
    AFK_NEED_OUT(2)
    if (!args_match(in, n_in, {AFK_T_IMAGE, AFK_T_INTEGER, AFK_T_INTEGER, AFK_T_INTEGER, AFK_T_BOOL}, out, n_out)) return AFK_ERR_PANIC;
    const int64_t range = in[3]->i;
    return argmap_op(in[0]->img, in[1]->i, in[2]->i, in[4]->b, false, -range, range, 2, out);
}

// ---- centrobaric wavelength (lib.rs:459-479, 1466-1582) ----------------------------------------------------------------------
AFK_OP(op_extract_centrobaric_wavelength) {
    AFK_NEED_OUT(1)
    if (!args_match(in, n_in, {AFK_T_IMAGE, AFK_T_INTEGER, AFK_T_INTEGER}, out, n_out)) return AFK_ERR_PANIC;
    const afk_image *img = in[0]->img;
    const int64_t start = in[1]->i, end = in[2]->i;
    Error e = range_checks(img, start, end);
    if (!e && !img->meta.has_meta) e = make_error(AFK_ERR_PANIC, "called `Option::unwrap()` on a `None` value");
    afk_image *res = nullptr;
    bool found = false;
    if (!e) {
        const int trypix = wave_axis(img, &found);
        Axis0Args a{};
        a.op = A0_CENTROID;
        a.start = start;
        a.end = end;
        a.use_wcs = wcs_axis(img, trypix, &a.crval, &a.crpix, &a.cdelt);
        e = op_axis0(img, a, nullptr, nullptr, &res, nullptr, "");
    }
    if (e) {
        set_err(&out[0], e);
        return e.status;
    }
    // &waveimg / &flux_sum (unit.rs:863-880): meta of waveimg (= the cube's), unit (u * waveunit) / u, tag dropped
    res->meta = img->meta;
    res->meta.has_tag = false;
    res->meta.tag.clear();
    bool h1;
    const std::string u1 = unit_mul(img->meta.has_unit, img->meta.unit, found, found ? "Ang" : "", &h1);
    res->meta.unit = unit_div(h1, u1, img->meta.has_unit, img->meta.unit, &res->meta.has_unit);
    res->meta.homogeneous = true;
    set_image(&out[0], res);
    return AFK_OK;
}

AFK_OP(op_extract_centrobaric_wavelength_with_mask) {
    AFK_NEED_OUT(1)
    if (!args_match(in, n_in, {AFK_T_IMAGE, AFK_T_IMAGE, AFK_T_IMAGE}, out, n_out)) return AFK_ERR_PANIC;
    const afk_image *img = in[0]->img, *lo = in[1]->img, *hi = in[2]->img;
    Error e = ok();
    // the reference indexes both masks with the cube's spatial index and panics on a mismatch
    auto mask_ok = [&](const afk_image *m) {
        if (img->ndim < 1 || m->ndim != img->ndim - 1) return false;
        for (int i = 0; i < m->ndim; ++i)
            if (m->shape[i] != img->shape[i + 1]) return false;
        return true;
    };
    if (!mask_ok(lo) || !mask_ok(hi)) e = make_error(AFK_ERR_PANIC, "ndarray: index out of bounds (mask shape differs from the image's spatial shape)");
    afk_image *res = nullptr;
    if (!e) {
        int dev = 0;
        cudaStream_t s;
        e = current_stream(&s, &dev);
        if (!e) e = same_device(lo, dev);
        if (!e) e = same_device(hi, dev);
    }
    if (!e) {
        Axis0Args a{};
        a.op = A0_CENTROID_MASK;
        a.start = 0;
        a.end = img->shape[0];
        a.use_wcs = wcs_axis(img, 2, &a.crval, &a.crpix, &a.cdelt); // pix2world(2, ..) (lib.rs:1562)
        e = op_axis0(img, a, lo, hi, &res, nullptr, "slice index out of range (start_mask > end_mask or end_mask > frame count)");
    }
    if (e) {
        set_err(&out[0], e);
        return e.status;
    }
    set_image(&out[0], res); // from_array(.., Unit::None)
    return AFK_OK;
}

// =============================================================================================================
// Registry
// =============================================================================================================
struct Entry {
    afk_transform_desc desc;
    std::vector<afk_input_slot> slots;
    std::vector<afk_type> outs;
};
static std::vector<Entry *> g_registry;
static std::once_flag g_registry_once;

struct SlotSpec {
    const char *name;
    afk_type type;
    afk_value *def; // nullptr: no default
};

static void add(const char *name, const char *kind, const char *description, int major, int minor, int patch,
                std::vector<SlotSpec> inputs, std::vector<afk_type> outputs, afk_plain_function f) {
    Entry *e = new Entry();
    for (auto &s : inputs) e->slots.push_back(afk_input_slot{s.name, s.type, s.def});
    e->outs = outputs;
    e->desc.name = name;
    e->desc.kind = kind;
    e->desc.description = description;
    e->desc.major = (uint8_t)major;
    e->desc.minor = (uint8_t)minor;
    e->desc.patch = (uint8_t)patch;
    e->desc.n_inputs = e->slots.size();
    e->desc.inputs = e->slots.data();
    e->desc.n_outputs = e->outs.size();
    e->desc.outputs = e->outs.data();
    e->desc.f = f;
    g_registry.push_back(e);
}

static afk_value *I(int64_t v) { return afk_value_integer(v); }
static afk_value *F(float v) { return afk_value_float(v); }
static afk_value *B(bool v) { return afk_value_bool(v); }
static afk_value *S(const char *v) { return afk_value_str(v); }

static void build_registry() {
    const afk_type IMG = AFK_T_IMAGE, INT = AFK_T_INTEGER, FLT = AFK_T_FLOAT, BOOL = AFK_T_BOOL, STR = AFK_T_STR, TC = AFK_T_TONECURVE;
    const afk_type PATHS = AFK_T_PATHS, FITS = AFK_T_FITS, F2 = AFK_T_FLOAT2, F3 = AFK_T_FLOAT3, F33 = AFK_T_FLOAT3X3, ROI = AFK_T_ROI,
                   MAP = AFK_T_MAP2DTO3DCOORDS;
    auto V3 = [](float a, float b, float c) { const float v[3] = {a, b, c}; return afk_value_float3(v); };
    // Names, kinds, versions, slot names, types and defaults follow lib.rs:141-747 (they are the RON on-disk contract,
    // aflak_cake/src/export.rs:80-85).  Descriptions are UI help text, reworded.
    add("open_fits", "01. Open File", "Open the n-th FITS file of a path list (headers are parsed, data stays on disk).", 1, 0, 0,
        {{"path", PATHS, nullptr}, {"n", INT, I(0)}}, {FITS}, op_open_fits);
    add("fits_to_image", "04. Extract part of data",
        "Decode the data unit of an HDU (by extension name, else by index) into a device-resident image.", 1, 0, 0,
        {{"fits", FITS, nullptr}, {"hdu_idx", INT, I(0)}, {"extension", STR, S("")}}, {IMG}, op_fits_to_image);
    add("make_float3", "02. Make new data", "Make a Float3 from 3 float values.", 1, 0, 0,
        {{"f1", FLT, F(0.0f)}, {"f2", FLT, F(0.0f)}, {"f3", FLT, F(0.0f)}}, {F3}, op_make_float3);
    add("make_plane3d", "02. Make new data",
        "2-D mesh of points p0 + i*dir1 + j*dir2 (0 <= i < count1, 0 <= j < count2) on a plane of the 3-D index space.", 1, 0, 0,
        {{"p0", F3, V3(0, 0, 0)}, {"dir1", F3, V3(0, 0, 1)}, {"dir2", F3, V3(0, 1, 0)}, {"count1", INT, I(1)}, {"count2", INT, I(1)}}, {MAP},
        op_make_plane3d);
    add("range_specification", "04. Extract part of data", "Keep frames start..end (end exclusive) of the 0th axis.", 0, 1, 0,
        {{"image", IMG, nullptr}, {"start", INT, I(0)}, {"end", INT, I(1)}}, {IMG}, op_range_specification);
    add("image_min_max", "04. Extract part of data", "Minimum (first output) and maximum (second output) of an image.", 1, 0, 0,
        {{"image", IMG, nullptr}}, {FLT, FLT}, op_image_min_max);
    add("ratio_from_bands", "02. Make new data",
        "Interpolation weights from band centres z1 < z < z2: off_ratio = 1 - (z - z1)/(z2 - z1) and its complement.", 1, 0, 0,
        {{"z", FLT, nullptr}, {"z1", FLT, nullptr}, {"z2", FLT, nullptr}}, {FLT, FLT}, op_ratio_from_bands);
    add("clip_image", "05. Convert data",
        "Replace values at or above the ceiling threshold and/or at or below the floor threshold with NaN.", 1, 0, 0,
        {{"image", IMG, nullptr}, {"ceiling_threshold", FLT, F(0.0f)}, {"ceiling", BOOL, B(false)}, {"floor_threshold", FLT, F(0.0f)}, {"floor", BOOL, B(false)}},
        {IMG}, op_clip_image);
    add("normalize_image", "05. Convert data", "Normalize image: (value - min) / (max - min).", 1, 0, 0,
        {{"image", IMG, nullptr}, {"min", FLT, F(0.0f)}, {"max", FLT, F(1.0f)}}, {IMG}, op_normalize_image);
    add("replace_nan_image", "05. Convert data", "Replace every NaN of the image with the placeholder value.", 1, 0, 0,
        {{"image", IMG, nullptr}, {"placeholder", FLT, F(0.0f)}}, {IMG}, op_replace_nan_image);
    add("convert_to_logscale", "05. Convert data",
        "Log-scale: y = log(a*x + 1) / log(a) with x = (value - v_min) / (v_max - v_min).", 1, 0, 0,
        {{"image", IMG, nullptr}, {"a", FLT, F(1000.0f)}, {"v_min", FLT, nullptr}, {"v_max", FLT, nullptr}}, {IMG}, op_convert_to_logscale);
    add("log10", "05. Convert data", "Base-10 logarithm of every value.", 1, 0, 0, {{"image", IMG, nullptr}}, {IMG}, op_log10);
    add("negation", "05. Convert data", "Negation: -image.", 1, 0, 0, {{"image", IMG, nullptr}}, {IMG}, op_negation);
    add("apply_tone_curve", "05. Convert data", "Map values through a tone-curve table.", 1, 0, 0,
        {{"image", IMG, nullptr}, {"tone_curve", TC, nullptr}}, {IMG}, op_apply_tone_curve);
    add("change_tag", "05. Convert data", "Set the visualization tag (supported: \"BPT\", \"color_image\").", 1, 0, 0,
        {{"image", IMG, nullptr}, {"tag", STR, S("")}}, {IMG}, op_change_tag);
    add("linear_composition", "06. Calculate", "Compose two images: a*u + b*v.", 1, 0, 0,
        {{"u", IMG, nullptr}, {"v", IMG, nullptr}, {"a", FLT, F(1.0f)}, {"b", FLT, F(1.0f)}}, {IMG}, op_linear_composition);
    add("image_multiplier", "06. Calculate", "Compose two images: u^a * v^b.", 1, 0, 0,
        {{"u", IMG, nullptr}, {"v", IMG, nullptr}, {"a", FLT, F(1.0f)}, {"b", FLT, F(1.0f)}}, {IMG}, op_image_multiplier);
    add("slice_one_frame", "07. Reduce dimension", "Take one frame of an n-dimensional dataset, giving an (n-1)-dimensional one.", 1, 0, 0,
        {{"image", IMG, nullptr}, {"frame", INT, I(0)}}, {IMG}, op_slice_one_frame);
    add("integral", "07. Reduce dimension",
        "Sum of frames start..end (end exclusive). Also outputs (start + end) / 2 and (end - start).", 1, 0, 0,
        {{"image", IMG, nullptr}, {"start", INT, I(0)}, {"end", INT, I(1)}}, {IMG, FLT, FLT}, op_integral);
    add("average", "07. Reduce dimension",
        "Mean of frames start..end (end exclusive). Also outputs (start + end) / 2 and (end - start).", 1, 0, 0,
        {{"image", IMG, nullptr}, {"start", INT, I(0)}, {"end", INT, I(1)}}, {IMG, FLT, FLT}, op_average);
    add("create_equivalent_width", "10. Create astronomy-specific map",
        "Equivalent width (i_off - i_on) * fl / i_off, sign flipped for emission; values above max become 0.", 0, 1, 0,
        {{"i_off", IMG, nullptr}, {"i_on", IMG, nullptr}, {"fl", FLT, F(1.0f)}, {"max", FLT, F(std::numeric_limits<float>::infinity())}, {"is_emission", BOOL, B(false)}},
        {IMG}, op_create_equivalent_width);
    add("create_emission_line_map", "10. Create astronomy-specific map",
        "Emission-line map (i_off - i_on) * fl, sign flipped for emission; values above max become 0.", 0, 1, 0,
        {{"i_off", IMG, nullptr}, {"i_on", IMG, nullptr}, {"fl", FLT, F(1.0f)}, {"max", FLT, F(std::numeric_limits<float>::infinity())}, {"is_emission", BOOL, B(false)}},
        {IMG}, op_create_emission_line_map);
    add("down_sampling", "02. Make new data", "Keep every n-th pixel in x and y (decimation).", 1, 0, 0,
        {{"image", IMG, nullptr}, {"n", INT, I(1)}}, {IMG}, op_down_sampling);
    add("variance", "07. Reduce dimension", "Sample variance (ddof = 1) of frames start..end (end exclusive).", 1, 0, 0,
        {{"image", IMG, nullptr}, {"start", INT, I(0)}, {"end", INT, I(1)}}, {IMG}, op_variance);
    add("stddev", "07. Reduce dimension", "Sample standard deviation (ddof = 1) of frames start..end (end exclusive).", 1, 0, 0,
        {{"image", IMG, nullptr}, {"start", INT, I(0)}, {"end", INT, I(1)}}, {IMG}, op_stddev);
    add("median", "07. Reduce dimension", "Per-pixel median of frames start..end (end exclusive).", 1, 0, 0,
        {{"image", IMG, nullptr}, {"start", INT, I(0)}, {"end", INT, I(1)}}, {IMG}, op_median);
    add("extract_argmin_max_wavelength", "08. Reduce dimension (index)",
        "Wavelength of the maximum (or minimum) flux of each pixel over frames start..end; second output is that flux.", 0, 1, 0,
        {{"image", IMG, nullptr}, {"start", INT, I(0)}, {"end", INT, I(1)}, {"is_min", BOOL, B(false)}}, {IMG, IMG},
        op_extract_argmin_max_wavelength);
    add("peak_based_wavelength_range", "08. Reduce dimension (index)",
        "Frame index of the flux peak of each pixel minus (first output) and plus (second output) range.", 0, 1, 0,
        {{"image", IMG, nullptr}, {"start", INT, I(0)}, {"end", INT, I(1)}, {"range", INT, I(1)}, {"is_min", BOOL, B(false)}}, {IMG, IMG},
        op_peak_based_wavelength_range);
    add("extract_centrobaric_wavelength", "08. Reduce dimension (index)",
        "Flux-weighted mean wavelength of each pixel over frames start..end: sum(f*w) / sum(f).", 0, 1, 0,
        {{"image", IMG, nullptr}, {"start", INT, I(0)}, {"end", INT, I(1)}}, {IMG}, op_extract_centrobaric_wavelength);
    add("extract_centrobaric_wavelength_with_mask", "08. Reduce dimension (index)",
        "Flux-weighted mean wavelength of each pixel over its own frame range [start_mask, end_mask).", 0, 1, 0,
        {{"image", IMG, nullptr}, {"start_mask", IMG, nullptr}, {"end_mask", IMG, nullptr}}, {IMG},
        op_extract_centrobaric_wavelength_with_mask);
    add("slice_3d_to_2d", "07. Reduce dimension", "Sample a 3-D dataset on the points of a 2-D mesh (nearest voxel).", 0, 1, 0,
        {{"image", IMG, nullptr}, {"map", MAP, nullptr}}, {IMG}, op_slice_3d_to_2d);
    add("extrude", "07. Reduce dimension", "Spectra of the pixels of a region of interest, one column per pixel, frames in reverse order.", 0, 1, 0,
        {{"image", IMG, nullptr}, {"roi", ROI, afk_value_default_for(AFK_T_ROI)}}, {IMG}, op_extrude);
    add("extract_wave", "07. Reduce dimension", "Sum of the region of interest in every frame: a 1-D spectrum.", 1, 0, 0,
        {{"image", IMG, nullptr}, {"roi", ROI, afk_value_default_for(AFK_T_ROI)}}, {IMG}, op_extract_wave);
    add("create_velocity_field_map", "10. Create astronomy-specific map", "Velocity v = c * (w - w_0) / w_0 with c = 3e5 km/s.", 1, 0, 0,
        {{"image", IMG, nullptr}, {"w_0", FLT, F(0.0f)}}, {IMG}, op_create_velocity_field_map);
    add("color_image_to_hsv", "09. Reduce dimension (color image)", "HSV channels of an RGB image.", 1, 0, 0, {{"image", IMG, nullptr}}, {IMG},
        op_color_image_to_hsv);
    add("color_image_from_rgb", "03. Make new data (color)", "Stack three 2-D images into one [3, ny, nx] colour image.", 1, 0, 0,
        {{"image_r", IMG, nullptr}, {"image_g", IMG, nullptr}, {"image_b", IMG, nullptr}}, {IMG}, op_color_image_from_rgb);
    add("histogram_transformation", "05. Convert data", "Midtone transfer function between the image minimum and maximum.", 0, 1, 0,
        {{"image", IMG, nullptr}, {"parameters", F3, V3(0.0f, 0.5f, 1.0f)}, {"_dynamic_range", F2, []() { const float v[2] = {0.0f, 1.0f}; return afk_value_float2(v); }()}},
        {IMG}, op_histogram_transformation);
    add("histogram_transformation_rgb", "05. Convert data", "Midtone transfer function per colour channel.", 0, 1, 0,
        {{"image", IMG, nullptr}, {"parameters", F33, []() { const float v[9] = {0.0f, 0.5f, 1.0f, 0.0f, 0.5f, 1.0f, 0.0f, 0.5f, 1.0f}; return afk_value_float3x3(v); }()},
         {"_dynamic_range", F2, []() { const float v[2] = {0.0f, 1.0f}; return afk_value_float2(v); }()}},
        {IMG}, op_histogram_transformation_rgb);
    add("apply_arcsinh_stretch", "05. Convert data", "Arcsinh stretch of a colour image: every channel scaled by asinh(l*beta) / (l*asinh(beta)).", 0, 1, 0,
        {{"image", IMG, nullptr}, {"beta", FLT, F(0.000000001f)}}, {IMG}, op_apply_arcsinh_stretch);
    // additive entry (not in the reference registry): the `&WcsArray / &WcsArray` operator of unit.rs:863-880
    add("ratio", "06. Calculate", "Ratio of two images: u / v.", 1, 0, 0, {{"u", IMG, nullptr}, {"v", IMG, nullptr}}, {IMG}, op_ratio);
}

void registry_init() { std::call_once(g_registry_once, build_registry); }

} // namespace afk

using namespace afk;

extern "C" {

size_t afk_transform_count(void) {
    registry_init();
    return g_registry.size();
}

const afk_transform_desc *afk_transform_at(size_t i) {
    registry_init();
    return i < g_registry.size() ? &g_registry[i]->desc : nullptr;
}

const afk_transform_desc *afk_get_transform(const char *name) {
    registry_init();
    if (!name) return nullptr;
    for (Entry *e : g_registry)
        if (strcmp(e->desc.name, name) == 0) return &e->desc;
    return nullptr;
}

afk_status afk_transform_call(const afk_transform_desc *t, const afk_value *const *inputs, size_t n_inputs, afk_result *outputs,
                              size_t n_outputs) {
    if (!t || !outputs) return fail(make_error(AFK_ERR_PANIC, "null transform or outputs"));
    for (size_t k = 0; k < n_outputs; ++k) outputs[k] = afk_result{AFK_OK, nullptr, nullptr};
    if (n_outputs != t->n_outputs) return fail(make_error(AFK_ERR_PANIC, "wrong number of output slots"));
    if (n_inputs < t->n_inputs) return fail(make_error(AFK_ERR_PANIC, "Missing input arguments!"));   // transform.rs:377-378
    if (n_inputs > t->n_inputs) return fail(make_error(AFK_ERR_PANIC, "Not all type consumed"));       // transform.rs:335-338
    std::vector<afk_value *> fed(t->n_inputs, nullptr);
    afk_status st = AFK_OK;
    for (size_t i = 0; i < t->n_inputs && st == AFK_OK; ++i) {
        const afk_value *v = inputs[i] ? inputs[i] : t->inputs[i].default_value;
        if (!v) {
            // dst/compute.rs:221-225 MissingDependency
            Error e = make_error(AFK_ERR_ARGUMENT, fmt("No output attached to input #%zu in '%s', cannot compute", i, t->name));
            for (size_t k = 0; k < n_outputs; ++k) { outputs[k].status = e.status; outputs[k].message = strdup(e.message.c_str()); }
            st = fail(e);
            break;
        }
        fed[i] = afk_value_convert(v, t->inputs[i].type_id); // feed (transform.rs:330-346)
        if (!fed[i]) {
            Error e = make_error(AFK_ERR_ARGUMENT, fmt("Cannot convert '%s' to '%s'", afk_type_name(v->type), afk_type_name(t->inputs[i].type_id)));
            for (size_t k = 0; k < n_outputs; ++k) { outputs[k].status = e.status; outputs[k].message = strdup(e.message.c_str()); }
            st = fail(e);
        }
    }
    if (st == AFK_OK) st = t->f(fed.data(), t->n_inputs, outputs, n_outputs);
    for (afk_value *v : fed) afk_value_release(v);
    return st;
}

// ---- fused paths -------------------------------------------------------------------------------------------------------------
afk_status afk_fused_average_ratio(const afk_image *cube, int64_t s0, int64_t e0, int64_t s1, int64_t e1, afk_image **out) {
    if (!cube || !out) return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "null argument"));
    const int64_t st[2] = {s0, s1}, en[2] = {e0, e1};
    for (int w = 0; w < 2; ++w) {
        Error e = try_into_unsigned("start", st[w]);
        if (!e) e = try_into_unsigned("end", en[w]);
        if (!e) e = is_sliceable_range("image", cube, st[w], en[w]);
        if (e) return fail(e);
    }
    afk_image *res = nullptr;
    if (Error e = op_reduce(cube, 2, st, en, EPI_MEAN_RATIO, nullptr, 1.0f, false, &res, nullptr)) return fail(e);
    // average -> make_slice; ratio -> meta of lhs, unit(u)/unit(u)
    meta_make_slice(cube, res);
    res->meta.unit = unit_div(cube->meta.has_unit, cube->meta.unit, cube->meta.has_unit, cube->meta.unit, &res->meta.has_unit);
    res->meta.homogeneous = true;
    *out = res;
    return AFK_OK;
}

afk_status afk_fused_logscale_minmax(const afk_image *img, float a, afk_image **out) {
    // convert_to_logscale(img, a, image_min_max(img).min, image_min_max(img).max) (src/templates.rs: the tail of the
    // equivalent-width template): the two floats never visit the host -- minmax_kernel leaves them on the device and the
    // logscale kernel reads them there.  Same arithmetic as the two operators (lib.rs:180-197, 1598-1611).
    if (!img || !out) return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "null argument"));
    cudaStream_t s;
    int dev;
    if (Error e = current_stream(&s, &dev)) return fail(e);
    if (Error e = same_device(img, dev)) return fail(e);
    afk_image *o = nullptr;
    if (Error e = image_new(img->ndim, img->shape, &o)) return fail(e);
    unsigned *mm = nullptr;
    const bool memo = img->range_valid && img->range; // the producer left the range beside the pixels: one launch, not three
    Error err = memo ? ok() : scratch_u32(s, 4, &mm);
    do {
        if (err) break;
        if ((err = image_wait(img, s))) break;
        cudaError_t ce = memo ? cudaSuccess : minmax_launch(img->ptr, img->len(), mm, sm_count(dev), s);
        if (ce == cudaSuccess) {
            MapArgs m{MAP_LOGSCALE, img->ptr, nullptr, a, 0.0f, 0.0f, 0, 0, img->len(), o->ptr};
            m.bc_minmax = memo ? img->range : mm;
            ce = map_launch(m, sm_count(dev), s);
        }
        if (ce != cudaSuccess) {
            err = cuda_error(ce, "fused logscale launch");
            break;
        }
        image_note_read(img, s);
        err = image_mark_ready(o, s);
    } while (0);
    if (mm) scratch_free(s, mm);
    if (err) {
        afk_image_release(o);
        return fail(err);
    }
    meta_clone(img, o); // convert_to_logscale clones its input (lib.rs:1598-1611)
    *out = o;
    return AFK_OK;
}

afk_status afk_fused_equivalent_width(const afk_image *cube, int64_t on_s, int64_t on_e, int64_t off1_s, int64_t off1_e, int64_t off2_s,
                                      int64_t off2_e, float fl, float max, int is_emission, afk_image **out, float minmax[2]) {
    if (!cube || !out) return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "null argument"));
    const int64_t st[3] = {on_s, off1_s, off2_s}, en[3] = {on_e, off1_e, off2_e};
    float mid[3];
    for (int w = 0; w < 3; ++w) {
        Error e = try_into_unsigned("start", st[w]);
        if (!e) e = try_into_unsigned("end", en[w]);
        if (!e) e = is_sliceable_range("image", cube, st[w], en[w]);
        if (e) return fail(e);
        mid[w] = ((float)st[w] + (float)en[w]) / 2.0f; // `middle` output of average (lib.rs:407)
    }
    // ratio_from_bands(z = mid_on, z1 = mid_off1, z2 = mid_off2) (lib.rs:226-246)
    const float z = mid[0], z1 = mid[1], z2 = mid[2];
    if (!(z1 < z && z < z2))
        return fail(unexpected(fmt("wrong magnitude correlation (%s < %s < %s)", rust_f32_display(z1).c_str(), rust_f32_display(z).c_str(),
                                   rust_f32_display(z2).c_str())));
    const float off_ratio_1 = (z - z1) / (z2 - z1);
    const float off_ratio_2 = 1.0f - off_ratio_1;
    // linear_composition(u = off1, v = off2, a = off_ratio_2, b = off_ratio_1) (templates.rs node 22)
    const float params[4] = {off_ratio_2, off_ratio_1, fl, max};
    afk_image *res = nullptr;
    if (Error e = op_reduce(cube, 3, st, en, EPI_EW, params, is_emission ? -1.0f : 1.0f, minmax != nullptr, &res, minmax)) return fail(e);
    ew_unit(cube, res);
    *out = res;
    return AFK_OK;
}

// ---- raw kernel layer --------------------------------------------------------------------------------------------------------------
afk_status afk_k_reduce(const float *cube, int64_t nz, int64_t n2, int nwin, const int64_t *starts, const int64_t *ends, int epilogue,
                        const float *params, int sum_mode, float *out) {
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
    // params are read for EPI_EW only: exactly 5 floats (a, b, fl, max, sign)
    if (epilogue == EPI_EW && !params) return fail(unexpected("the equivalent-width epilogue needs params[5] = {a, b, fl, max, sign}"));
    for (int i = 0; i < 4; ++i) a.p[i] = epilogue == EPI_EW ? params[i] : 0.0f;
    a.ew_sign = epilogue == EPI_EW ? params[4] : 1.0f;
    a.out = out;
    a.minmax = nullptr;
    if (n2 <= 0) return AFK_OK;
    cudaStream_t s;
    int dev;
    if (Error e = current_stream(&s, &dev)) return fail(e);
    ReducePlan plan = reduce_plan(a, sm_count(dev));
    ReduceWorkspace ws;
    if (Error e = reduce_workspace(plan, dev, s, &ws)) return fail(e);
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
    if (ce != cudaSuccess) return fail(cuda_error(ce, "reduce_kernel launch"));
    return AFK_OK;
}

// Which instantiation and launch shape afk_k_reduce / afk_gather_reduce use for this problem, as text
// ("reduce_kernel<VEC=4,COMP=1,NW=2,U=8,MINB=3,KG=1> grid=444 block=256 ..."), so that reports name the kernel that ran.
int afk_k_reduce_describe(const float *cube, int64_t n2, int nwin, const int64_t *starts, const int64_t *ends, int sum_mode,
                          const float *out, char *buf, size_t cap) {
    if (!starts || !ends || !buf || cap == 0 || nwin < 1 || nwin > kMaxWindows) return fail(unexpected("bad arguments to afk_k_reduce_describe"));
    ReduceArgs a{};
    a.cube = cube;
    a.n2 = n2;
    a.nwin = nwin;
    for (int w = 0; w < nwin; ++w) {
        a.start[w] = starts[w];
        a.end[w] = ends[w];
    }
    a.mode = sum_mode;
    a.out = const_cast<float *>(out);
    cudaStream_t s;
    int dev;
    if (Error e = current_stream(&s, &dev)) return fail(e);
    const ReducePlan pl = reduce_plan(a, sm_count(dev));
    reduce_describe(a, pl, buf, cap);
    return AFK_OK;
}

} // extern "C"
