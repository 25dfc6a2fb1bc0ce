// ops_spatial.cpp -- host side of the remaining elementwise / gather operators: make_float3, make_plane3d,
// slice_3d_to_2d, extrude, extract_wave, create_velocity_field_map, histogram_transformation(+_rgb),
// apply_arcsinh_stretch, color_image_to_hsv, color_image_from_rgb.
//
// Reference: aflak_primitives/src/lib.rs:198-225, 360-383, 509-631 (registry); :928-1169, :1233-1256, :1795-1856,
// :1873-1892 (run_* bodies); roi.rs:12-37; precond.rs:59-74 (dim_is!).
#include <cfloat>
#include <cmath>
#include <cstring>
#include <limits>

#include "host.h"

namespace afk {

#define AFK_OPX(name) afk_status name(const afk_value *const *in, size_t n_in, afk_result *out, size_t n_out)
#define AFK_NEED_OUT(n)                                                                            \
    if (n_out != (n)) return fail(make_error(AFK_ERR_PANIC, "wrong number of output slots"));

static afk_status finish(afk_result *out, const Error &e, afk_image *res) {
    if (e) {
        if (res) afk_image_release(res);
        set_err(&out[0], e);
        return e.status;
    }
    set_image(&out[0], res);
    return AFK_OK;
}

// dim_is!(image, n) (precond.rs:59-74)
static Error dim_is(const char *name, const afk_image *img, int expected) {
    if (img->ndim == expected) return ok();
    return unexpected(fmt("'%s' is a %d-dimensional image, while dimension %d was expected", name, img->ndim, expected));
}

struct OpCtx {
    cudaStream_t s = nullptr;
    int dev = 0;
};
static Error ctx_for(OpCtx *c, std::initializer_list<const afk_image *> imgs) {
    if (Error e = current_stream(&c->s, &c->dev)) return e;
    for (const afk_image *i : imgs) {
        if (Error e = same_device(i, c->dev)) return e;
        if (Error e = image_wait(i, c->s)) return e;
    }
    return ok();
}

// ---- make_float3 / make_plane3d (lib.rs:198-225, 1130-1145, 1298-1300): host values -----------------------------------
AFK_OPX(op_make_float3) {
    AFK_NEED_OUT(1)
    if (!args_match(in, n_in, {AFK_T_FLOAT, AFK_T_FLOAT, AFK_T_FLOAT}, out, n_out)) return AFK_ERR_PANIC;
    const float v[3] = {in[0]->f[0], in[1]->f[0], in[2]->f[0]};
    set_ok(&out[0], afk_value_float3(v));
    return AFK_OK;
}

AFK_OPX(op_make_plane3d) {
    AFK_NEED_OUT(1)
    if (!args_match(in, n_in, {AFK_T_FLOAT3, AFK_T_FLOAT3, AFK_T_FLOAT3, AFK_T_INTEGER, AFK_T_INTEGER}, out, n_out)) return AFK_ERR_PANIC;
    const float *p0 = in[0]->f, *d1 = in[1]->f, *d2 = in[2]->f;
    const int64_t count1 = in[3]->i, count2 = in[4]->i;
    Error e = try_into_unsigned("count1", count1);
    if (!e) e = try_into_unsigned("count2", count2);
    if (!e && count1 * count2 > ((int64_t)1 << 31)) e = unexpected("make_plane3d: mesh too large");
    if (e) {
        set_err(&out[0], e);
        return e.status;
    }
    // Array2::from_shape_fn((count2, count1), |(j, i)| [x0 + i*dx1 + j*dx2, ...]) in f32, left to right
    std::vector<float> map((size_t)(3 * count1 * count2));
    for (int64_t j = 0; j < count2; ++j)
        for (int64_t i = 0; i < count1; ++i) {
            const float fi = (float)i, fj = (float)j;
            float *m = &map[(size_t)(3 * (j * count1 + i))];
            for (int c = 0; c < 3; ++c) m[c] = (p0[c] + fi * d1[c]) + fj * d2[c];
        }
    set_ok(&out[0], afk_value_map2dto3d(map.data(), count2, count1));
    return AFK_OK;
}

// ---- slice_3d_to_2d (lib.rs:360-367, 928-1128) ----------------------------------------------------------------------------
// MapReverseParams: constant per-component step of the mesh along its two directions, or None when it varies
struct DirValue {
    int state = 0; // 0 unset, 1 some, 2 none
    float v = 0.0f;
    void update(float t) {
        if (state == 0) { state = 1; v = t; }
        else if (state == 1 && fabsf(v - t) > FLT_EPSILON) state = 2;
    }
    bool some() const { return state == 1; }
};

// sliced_axes (lib.rs:1020-1075): Some([(axis, start, step); 2]) when the mesh is aligned with two axes
static bool sliced_axes(const std::vector<float> &map, int64_t n, int64_t m, int axis[2], float start[2], float step[2]) {
    DirValue d1[3], d2[3];
    auto at = [&](int64_t i, int64_t j, int c) { return map[(size_t)(3 * (i * m + j) + c)]; };
    for (int64_t i = 0; i < n; ++i)
        for (int64_t j = 0; j + 1 < m; ++j)
            for (int c = 0; c < 3; ++c) d1[c].update(at(i, j + 1, c) - at(i, j, c));
    for (int64_t i = 0; i + 1 < n; ++i)
        for (int64_t j = 0; j < m; ++j)
            for (int c = 0; c < 3; ++c) d2[c].update(at(i + 1, j, c) - at(i, j, c));
    const float EPS = FLT_EPSILON;
    bool remove[3];
    for (int c = 0; c < 3; ++c) {
        remove[c] = false;
        if (d1[c].some() && d2[c].some()) {
            const float a = d1[c].v, b = d2[c].v;
            if ((a < EPS && b > EPS) || (a > EPS && b < EPS)) remove[c] = false; // KeepDirection
            else if (a < EPS && b < EPS) remove[c] = true;                       // RemoveDirection
        }
    }
    if (n * m == 0) return false; // origin = map.get((0, 0)) is None
    auto factor = [&](int c, float *f) { // non_zero_factor
        if (!(d1[c].some() && d2[c].some())) return false;
        *f = fabsf(d1[c].v) < EPS ? d2[c].v : d1[c].v;
        return true;
    };
    const float origin[3] = {map[0], map[1], map[2]};
    float fa, fb;
    if (remove[0] && !remove[1] && !remove[2]) {
        if (!factor(2, &fa) || !factor(1, &fb)) return false;
        axis[0] = 0; start[0] = origin[2]; step[0] = fa;
        axis[1] = 1; start[1] = origin[1]; step[1] = fb;
        return true;
    }
    if (!remove[0] && remove[1] && !remove[2]) {
        if (!factor(2, &fa) || !factor(0, &fb)) return false;
        axis[0] = 0; start[0] = origin[2]; step[0] = fa;
        axis[1] = 2; start[1] = origin[0]; step[1] = fb;
        return true;
    }
    if (!remove[0] && !remove[1] && remove[2]) {
        if (!factor(1, &fa) || !factor(0, &fb)) return false;
        axis[0] = 1; start[0] = origin[1]; step[0] = fa;
        axis[1] = 2; start[1] = origin[0]; step[1] = fb;
        return true;
    }
    return false;
}

// make_slice(&[(i, start, step)...]) on the linear WCS: the kept FITS axes are re-based so that new pixel p' sits at old
// pixel start + step * p' (wcs.slice(indices) then wcs.transform(i, start, step), unit.rs:424-451); other axes reset
static void meta_make_slice_axes(const afk_image *src, afk_image *dst, int n_axes, const int *axis, const float *start, const float *step) {
    dst->meta = ImageMeta();
    dst->meta.has_unit = src->meta.has_unit;
    dst->meta.unit = src->meta.unit;
    dst->meta.homogeneous = src->meta.homogeneous;
    dst->meta.has_meta = src->meta.has_meta;
    if (!src->meta.has_meta) return;
    if (src->meta.wcs.size() == 12) {
        dst->meta.wcs.assign(12, 0.0);
        for (int a = 0; a < 4; ++a) dst->meta.wcs[3 * a + 1] = dst->meta.wcs[3 * a + 2] = 1.0;
    }
    for (int k = 0; k < n_axes; ++k) {
        const int a = axis[k];
        dst->meta.axes[a] = src->meta.axes[a]; // axes[*i] = meta.axes[*i].clone()
        if (src->meta.wcs.size() == 12) {
            const double crval = src->meta.wcs[3 * a], crpix = src->meta.wcs[3 * a + 1], cdelt = src->meta.wcs[3 * a + 2];
            dst->meta.wcs[3 * a + 0] = crval + cdelt * ((double)start[k] + (1.0 - crpix));
            dst->meta.wcs[3 * a + 1] = 1.0;
            dst->meta.wcs[3 * a + 2] = cdelt * (double)step[k];
        }
    }
}

AFK_OPX(op_slice_3d_to_2d) {
    AFK_NEED_OUT(1)
    if (!args_match(in, n_in, {AFK_T_IMAGE, AFK_T_MAP2DTO3DCOORDS}, out, n_out)) return AFK_ERR_PANIC;
    const afk_image *img = in[0]->img;
    const afk_value *mv = in[1];
    Error e = dim_is("image", img, 3);
    OpCtx c;
    if (!e) e = ctx_for(&c, {img});
    afk_image *res = nullptr;
    float *dmap = nullptr;
    unsigned long long *bad = nullptr;
    const int64_t n = mv->map_rows * mv->map_cols;
    do {
        if (e) break;
        const int64_t shape[2] = {mv->map_rows, mv->map_cols};
        if ((e = image_new(2, shape, &res))) break;
        if (n == 0) {
            e = image_mark_ready(res, c.s);
            break;
        }
        // the mesh is an immutable host value: its device copy is made once and kept with the value
        cudaError_t ce = cudaSuccess;
        {
            afk_value *mut = const_cast<afk_value *>(mv);
            std::lock_guard<std::mutex> lk(mut->map_mu);
            if (mut->map_dev && mut->map_dev_device != c.dev) {
                cudaFree(mut->map_dev);
                mut->map_dev = nullptr;
            }
            if (!mut->map_dev) {
                ce = cudaMalloc(reinterpret_cast<void **>(&mut->map_dev), (size_t)n * 3 * sizeof(float));
                if (ce == cudaSuccess) ce = cudaMemcpy(mut->map_dev, mv->map.data(), (size_t)n * 3 * sizeof(float), cudaMemcpyHostToDevice);
                mut->map_dev_device = c.dev;
            }
            dmap = mut->map_dev;
        }
        if (ce == cudaSuccess) ce = cudaMallocAsync(reinterpret_cast<void **>(&bad), sizeof *bad, c.s);
        if (ce == cudaSuccess) ce = cudaMemsetAsync(bad, 0xFF, sizeof *bad, c.s);
        if (ce == cudaSuccess) ce = slice3d_launch(img->ptr, img->shape[0], img->shape[1], img->shape[2], dmap, n, res->ptr, bad, sm_count(c.dev), c.s);
        unsigned long long hbad = 0;
        if (ce == cudaSuccess) ce = cudaMemcpyAsync(&hbad, bad, sizeof hbad, cudaMemcpyDeviceToHost, c.s);
        if (ce == cudaSuccess) ce = cudaStreamSynchronize(c.s); // the bounds check needs the flag, and the map is host memory
        if (ce != cudaSuccess) {
            e = cuda_error(ce, "slice3d_kernel");
            break;
        }
        image_note_read(img, c.s);
        if (hbad != ~0ull) {
            const float *p = &mv->map[(size_t)(3 * hbad)];
            e = unexpected(fmt("Input maps to out of bound pixel!: [%s, %s, %s]", rust_f32_display(p[0]).c_str(), rust_f32_display(p[1]).c_str(),
                               rust_f32_display(p[2]).c_str()));
            break;
        }
        e = image_mark_ready(res, c.s);
    } while (0);
    if (bad) cudaFreeAsync(bad, c.s);
    if (!e) {
        afk_value *mut = const_cast<afk_value *>(mv);
        {
            std::lock_guard<std::mutex> lk(mut->map_mu);
            if (mut->map_sliced < 0)
                mut->map_sliced = sliced_axes(mv->map, mv->map_rows, mv->map_cols, mut->map_axis, mut->map_start, mut->map_step) ? 1 : 0;
        }
        if (mut->map_sliced == 1) meta_make_slice_axes(img, res, 2, mut->map_axis, mut->map_start, mut->map_step);
        else { // WcsArray::from_array(array): no meta, unit kept (array = image.array().with_new_value(..))
            res->meta = ImageMeta();
            res->meta.has_unit = img->meta.has_unit;
            res->meta.unit = img->meta.unit;
            res->meta.homogeneous = img->meta.homogeneous;
        }
    }
    return finish(out, e, res);
}

// ---- regions of interest (roi.rs:12-37) ---------------------------------------------------------------------------------
// ROI::filter on a [ny, nx] frame -> linear offsets y * nx + x in the reference's iteration order.
//   All:       for j in 0..nx { for i in 0..ny { data.get([j, i]).unwrap() } }: row j, column i -- the reference's loop
//              bounds are swapped, so it only works (and is plain row-major order) on square frames and panics otherwise
//   PixelList: (i, j) -> data.get([j, i]); pixels outside the frame are skipped
static Error roi_offsets(const afk_value *roi, int64_t ny, int64_t nx, bool *whole_frame, std::vector<int64_t> *offsets) {
    *whole_frame = false;
    offsets->clear();
    if (roi->roi_all) {
        if (ny != nx) return make_error(AFK_ERR_PANIC, "called `Option::unwrap()` on a `None` value (ROI::All on a non-square frame)");
        *whole_frame = true;
        return ok();
    }
    for (size_t p = 0; p + 1 < roi->roi_xy.size(); p += 2) {
        const uint64_t i = roi->roi_xy[p], j = roi->roi_xy[p + 1];
        if (j < (uint64_t)ny && i < (uint64_t)nx) offsets->push_back((int64_t)j * nx + (int64_t)i);
    }
    return ok();
}

// Device copy of the offsets, built once per (Roi value, frame shape, device) and kept with the immutable value like the
// mesh of slice_3d_to_2d: re-evaluating a node whose region did not change costs no upload and no synchronisation.
static Error roi_device_offsets(const afk_value *roi, int64_t ny, int64_t nx, int dev, cudaStream_t s, bool *whole_frame, const int64_t **d,
                                int64_t *count) {
    *d = nullptr;
    *count = 0;
    *whole_frame = false;
    if (roi->roi_all) {
        std::vector<int64_t> none;
        Error e = roi_offsets(roi, ny, nx, whole_frame, &none);
        if (!e) *count = ny * nx;
        return e;
    }
    afk_value *mut = const_cast<afk_value *>(roi);
    std::lock_guard<std::mutex> lk(mut->map_mu);
    if (mut->roi_dev_device != dev || mut->roi_ny != ny || mut->roi_nx != nx) {
        if (mut->roi_dev) cudaFree(mut->roi_dev);
        mut->roi_dev = nullptr;
        mut->roi_dev_device = -1;
        std::vector<int64_t> off;
        Error e = roi_offsets(roi, ny, nx, whole_frame, &off);
        if (e) return e;
        if (!off.empty()) {
            // from the stream-ordered pool (a plain cudaMalloc costs more than the kernels that follow); `off` is pageable
            AFK_CUDA(cudaMallocAsync(reinterpret_cast<void **>(&mut->roi_dev), off.size() * sizeof(int64_t), s));
            AFK_CUDA(cudaMemcpyAsync(mut->roi_dev, off.data(), off.size() * sizeof(int64_t), cudaMemcpyHostToDevice, s));
            AFK_CUDA(cudaStreamSynchronize(s));
        }
        mut->roi_count = (int64_t)off.size();
        mut->roi_ny = ny;
        mut->roi_nx = nx;
        mut->roi_dev_device = dev;
    }
    *d = mut->roi_dev;
    *count = mut->roi_count;
    return ok();
}

AFK_OPX(op_extract_wave) { // lib.rs:376-383, 1147-1169
    AFK_NEED_OUT(1)
    if (!args_match(in, n_in, {AFK_T_IMAGE, AFK_T_ROI}, out, n_out)) return AFK_ERR_PANIC;
    const afk_image *img = in[0]->img;
    Error e = dim_is("image", img, 3);
    OpCtx c;
    if (!e) e = ctx_for(&c, {img});
    afk_image *res = nullptr;
    const int64_t *doff = nullptr;
    void *partials = nullptr;
    do {
        if (e) break;
        const int64_t nz = img->shape[0], ny = img->shape[1], nx = img->shape[2];
        bool whole;
        int64_t count;
        if ((e = roi_device_offsets(in[1], ny, nx, c.dev, c.s, &whole, &doff, &count))) break;
        if ((e = image_new(1, &nz, &res))) break;
        const int mode = effective_sum_mode();
        cudaError_t ce = cudaSuccess;
        if (count == 0 || (!whole && !doff)) {
            ce = cudaMemsetAsync(res->ptr, 0, res->bytes, c.s); // res = 0.0 for every frame
        } else {
            if (mode != SUM_SEQUENTIAL) {
                int chunks;
                int64_t chunk;
                const size_t pb = wave_partial_bytes(nz, count, &chunks, &chunk);
                ce = cudaMallocAsync(&partials, pb ? pb : 16, c.s);
            }
            if (ce == cudaSuccess) ce = wave_launch(img->ptr, nz, ny * nx, whole ? nullptr : doff, count, mode, partials, res->ptr, c.s);
        }
        if (ce != cudaSuccess) {
            e = cuda_error(ce, "wave kernels");
            break;
        }
        image_note_read(img, c.s);
        e = image_mark_ready(res, c.s);
    } while (0);
    if (partials) cudaFreeAsync(partials, c.s);
    if (!e) {
        // image.make_slice(&[(2, 0.0, 1.0)], ..): FITS axis 2 (the spectral axis) is the only one kept; unit kept, tag dropped
        const int axis[1] = {2};
        const float start[1] = {0.0f}, step[1] = {1.0f};
        meta_make_slice_axes(img, res, 1, axis, start, step);
    }
    return finish(out, e, res);
}

AFK_OPX(op_extrude) { // lib.rs:368-375, 1233-1256
    AFK_NEED_OUT(1)
    if (!args_match(in, n_in, {AFK_T_IMAGE, AFK_T_ROI}, out, n_out)) return AFK_ERR_PANIC;
    const afk_image *img = in[0]->img;
    Error e = dim_is("image", img, 3);
    OpCtx c;
    if (!e) e = ctx_for(&c, {img});
    afk_image *res = nullptr;
    const int64_t *doff = nullptr;
    do {
        if (e) break;
        const int64_t nz = img->shape[0], ny = img->shape[1], nx = img->shape[2];
        bool whole;
        int64_t pushed;
        if ((e = roi_device_offsets(in[1], ny, nx, c.dev, c.s, &whole, &doff, &pushed))) break;
        // new_size = (wave_size, roi.datalen()); ROI::All has datalen 0, and skipped pixels leave the vector short:
        // Array::from_shape_vec(..).unwrap() panics unless the lengths agree
        const int64_t datalen = in[1]->roi_all ? 0 : (int64_t)(in[1]->roi_xy.size() / 2);
        if (nz * pushed != nz * datalen) {
            e = make_error(AFK_ERR_PANIC, "called `Result::unwrap()` on an `Err` value: ShapeError/OutOfBounds: out of bounds indexing");
            break;
        }
        const int64_t shape[2] = {nz, datalen};
        if ((e = image_new(2, shape, &res))) break;
        cudaError_t ce = extrude_launch(img->ptr, nz, ny * nx, doff, datalen, res->ptr, sm_count(c.dev), c.s);
        if (ce != cudaSuccess) {
            e = cuda_error(ce, "extrude_kernel");
            break;
        }
        image_note_read(img, c.s);
        e = image_mark_ready(res, c.s);
    } while (0);
    if (!e) { // from_array(Dimensioned::new(waveimg, unit)): unit kept, nothing else
        res->meta = ImageMeta();
        res->meta.has_unit = img->meta.has_unit;
        res->meta.unit = img->meta.unit;
    }
    return finish(out, e, res);
}

// ---- create_velocity_field_map (lib.rs:509-526) -----------------------------------------------------------------------------
AFK_OPX(op_create_velocity_field_map) {
    AFK_NEED_OUT(1)
    if (!args_match(in, n_in, {AFK_T_IMAGE, AFK_T_FLOAT}, out, n_out)) return AFK_ERR_PANIC;
    const afk_image *img = in[0]->img;
    afk_image *res = nullptr;
    Error e = op_map(MAP_VELOCITY, img, nullptr, in[1]->f[0], 0, 0, 0, 0, &res);
    if (!e) {
        // (image.array() - w_0[Ang]) * c[km/s] / w_0[Ang]; meta and tag of the input are re-attached
        res->meta = img->meta;
        bool h1, h2;
        const std::string u1 = unit_mul(img->meta.has_unit, img->meta.unit, true, "km/s", &h1);
        res->meta.unit = unit_div(h1, u1, true, "Ang", &h2);
        res->meta.has_unit = h2;
        res->meta.homogeneous = img->meta.homogeneous && img->meta.has_unit && img->meta.unit == "Ang";
    }
    return finish(out, e, res);
}

// ---- histogram_transformation (lib.rs:562-591) and _rgb (lib.rs:592-631) -----------------------------------------------
// min / max as in the operators' own loops: start at f32::MAX / f32::MIN, Rust min/max skip a NaN operand
static Error hist_min_max(const afk_image *view, float *mn, float *mx) {
    bool any;
    float a = 0, b = 0;
    if (view->len() == 0) {
        *mn = FLT_MAX;
        *mx = -FLT_MAX;
        return ok();
    }
    if (Error e = op_min_max(view, &a, &b, &any)) return e;
    *mn = any ? fminf(FLT_MAX, a) : FLT_MAX;
    *mx = any ? fmaxf(-FLT_MAX, b) : -FLT_MAX;
    return ok();
}

static Error hist_trans_launch(const float *src, float *dst, int64_t n, float mn, float mx, const float p[3], const OpCtx &c) {
    MapArgs m{MAP_HISTTRANS, src, nullptr, mn, mx, p[0], 0, 0, n, dst, p[1], p[2]};
    const cudaError_t ce = map_launch(m, sm_count(c.dev), c.s);
    return ce == cudaSuccess ? ok() : cuda_error(ce, "map_kernel launch");
}

AFK_OPX(op_histogram_transformation) {
    AFK_NEED_OUT(1)
    if (!args_match(in, n_in, {AFK_T_IMAGE, AFK_T_FLOAT3, AFK_T_FLOAT2}, out, n_out)) return AFK_ERR_PANIC;
    const afk_image *img = in[0]->img;
    OpCtx c;
    Error e = ctx_for(&c, {img});
    float mn = 0, mx = 0;
    if (!e) e = hist_min_max(img, &mn, &mx);
    afk_image *res = nullptr;
    if (!e) e = image_new(img->ndim, img->shape, &res);
    if (!e) e = hist_trans_launch(img->ptr, res->ptr, img->len(), mn, mx, in[1]->f, c);
    if (!e) {
        image_note_read(img, c.s);
        e = image_mark_ready(res, c.s);
    }
    if (!e) res->meta = img->meta; // image.clone(): meta, unit and tag kept
    return finish(out, e, res);
}

AFK_OPX(op_histogram_transformation_rgb) {
    AFK_NEED_OUT(1)
    if (!args_match(in, n_in, {AFK_T_IMAGE, AFK_T_FLOAT3X3, AFK_T_FLOAT2}, out, n_out)) return AFK_ERR_PANIC;
    const afk_image *img = in[0]->img;
    Error e = ok();
    // dim[0], dim[1], dim[2] index the shape and parameters[c] the 3x3 table: anything else is an index panic
    if (img->ndim != 3) e = make_error(AFK_ERR_PANIC, fmt("index out of bounds: the image has %d dimensions, 3 are indexed", img->ndim));
    else if (img->shape[0] > 3 && img->shape[1] * img->shape[2] > 0) e = make_error(AFK_ERR_PANIC, "index out of bounds: the len is 3 but the index is 3");
    OpCtx c;
    if (!e) e = ctx_for(&c, {img});
    afk_image *res = nullptr;
    if (!e) e = image_new(3, img->shape, &res);
    const int64_t n = e ? 0 : img->shape[1] * img->shape[2];
    for (int64_t ch = 0; !e && ch < img->shape[0]; ++ch) {
        afk_image view; // stack view of one channel for the min/max pass
        view.ptr = img->ptr + ch * n;
        view.owned = false;
        view.device = img->device;
        view.ndim = 2;
        view.shape[0] = img->shape[1];
        view.shape[1] = img->shape[2];
        view.bytes = (size_t)n * sizeof(float);
        view.rc.store(1 << 20); // never released
        float mn, mx;
        e = hist_min_max(&view, &mn, &mx);
        if (!e) e = hist_trans_launch(view.ptr, res->ptr + ch * n, n, mn, mx, in[1]->f + 3 * ch, c);
    }
    if (!e) {
        image_note_read(img, c.s);
        e = image_mark_ready(res, c.s);
    }
    if (!e) { // from_array_and_tag(Dimensioned::new(img, Unit::None), tag)
        res->meta = ImageMeta();
        res->meta.has_tag = img->meta.has_tag;
        res->meta.tag = img->meta.tag;
    }
    return finish(out, e, res);
}

// ---- colour-image operators (lib.rs:317-324, 538-553, 1795-1856, 1873-1892) ---------------------------------------------
static afk_status color3_op(int op, const afk_image *img, float beta, bool keep_tag, afk_result *out) {
    Error e = dim_is("image", img, 3);
    // data[0], data[1], data[2] of every pixel: fewer than three channels is an index panic (no pixels: nothing is indexed)
    if (!e && img->shape[0] < 3 && img->shape[1] * img->shape[2] > 0) e = make_error(AFK_ERR_PANIC, fmt("ndarray: index 2 is out of bounds for array of shape [%lld]", (long long)img->shape[0]));
    OpCtx c;
    if (!e) e = ctx_for(&c, {img});
    afk_image *res = nullptr;
    if (!e) e = image_new(3, img->shape, &res);
    if (!e) {
        const int64_t n = img->shape[1] * img->shape[2];
        cudaError_t ce = cudaSuccess;
        // `out = image.clone()`: channels beyond the third are copied unchanged
        if (img->shape[0] > 3) ce = cudaMemcpyAsync(res->ptr + 3 * n, img->ptr + 3 * n, (size_t)(img->shape[0] - 3) * n * sizeof(float), cudaMemcpyDeviceToDevice, c.s);
        if (ce == cudaSuccess) ce = color3_launch(op, img->ptr, n, beta, res->ptr, sm_count(c.dev), c.s);
        if (ce != cudaSuccess) e = cuda_error(ce, "color3_kernel");
    }
    if (!e) {
        image_note_read(img, c.s);
        e = image_mark_ready(res, c.s);
    }
    if (!e) {
        res->meta = ImageMeta(); // Unit::None, no meta
        if (keep_tag) {
            res->meta.has_tag = img->meta.has_tag;
            res->meta.tag = img->meta.tag;
        }
    }
    return finish(out, e, res);
}

AFK_OPX(op_apply_arcsinh_stretch) {
    AFK_NEED_OUT(1)
    if (!args_match(in, n_in, {AFK_T_IMAGE, AFK_T_FLOAT}, out, n_out)) return AFK_ERR_PANIC;
    return color3_op(C3_ARCSINH, in[0]->img, in[1]->f[0], true, out);
}

AFK_OPX(op_color_image_to_hsv) {
    AFK_NEED_OUT(1)
    if (!args_match(in, n_in, {AFK_T_IMAGE}, out, n_out)) return AFK_ERR_PANIC;
    return color3_op(C3_TO_HSV, in[0]->img, 0.0f, false, out);
}

AFK_OPX(op_color_image_from_rgb) {
    AFK_NEED_OUT(1)
    if (!args_match(in, n_in, {AFK_T_IMAGE, AFK_T_IMAGE, AFK_T_IMAGE}, out, n_out)) return AFK_ERR_PANIC;
    const afk_image *r = in[0]->img, *g = in[1]->img, *b = in[2]->img;
    Error e = dim_is("image_r", r, 2);
    if (!e) e = dim_is("image_g", g, 2);
    if (!e) e = dim_is("image_b", b, 2);
    if (!e) e = are_same_dim("image_r", r, "image_b", b);
    if (!e) e = are_same_dim("image_b", b, "image_g", g);
    OpCtx c;
    if (!e) e = ctx_for(&c, {r, g, b});
    afk_image *res = nullptr;
    if (!e) {
        const int64_t shape[3] = {3, r->shape[0], r->shape[1]};
        e = image_new(3, shape, &res);
    }
    if (!e) {
        const afk_image *src[3] = {r, g, b};
        for (int ch = 0; ch < 3 && !e; ++ch) {
            if (src[ch]->bytes == 0) continue;
            const cudaError_t ce = cudaMemcpyAsync(res->ptr + (size_t)ch * r->len(), src[ch]->ptr, src[ch]->bytes, cudaMemcpyDeviceToDevice, c.s);
            if (ce != cudaSuccess) e = cuda_error(ce, "cudaMemcpyAsync(D2D)");
            image_note_read(src[ch], c.s);
        }
    }
    if (!e) e = image_mark_ready(res, c.s);
    return finish(out, e, res); // from_array(.., Unit::None): a fresh ImageMeta
}

} // namespace afk
