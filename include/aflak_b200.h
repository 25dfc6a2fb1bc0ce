/*
 * aflak_b200.h -- C ABI of the B200-native replacement for aflak's data-parallel hot path.
 *
 * This is the drop-in boundary: the entry points below are what a Rust `-sys` crate inside
 * aflak_primitives would bind.  Plain pointers and sizes only.
 * Reference citations are relative to /root/reference/src/.
 *
 *   value model      mirrors `enum IOValue`               aflak_primitives/src/lib.rs:66-83
 *   errors           mirror  `enum IOErr`                 aflak_primitives/src/lib.rs:109-130
 *   registry         mirrors `TRANSFORMATIONS` + `cake::Transform` / `Algorithm::Function`
 *                                                         aflak_primitives/src/lib.rs:141-758, aflak_cake/src/transform.rs:33-78
 *   call             mirrors `TransformCaller::{feed,call}` aflak_cake/src/transform.rs:324-397
 *   display          mirrors `ColorLUT` + `make_raw_image` aflak_plot/src/imshow/lut.rs, image.rs:20-84
 *   evaluator+cache  mirrors `DST::compute` + `Cache`     aflak_cake/src/dst/compute.rs:153-274, cache.rs:18-148
 *
 * Threading: every entry point is re-entrant.  Each calling thread gets its own CUDA stream per device; image
 * handles carry a ready-event so chained operators on different threads never need a host sync.
 * Errors never abort: reference panics ("Unexpected argument!", index out of bounds) come back as AFK_ERR_PANIC.
 * There is no CPU fallback: without a CUDA device every compute entry point returns AFK_ERR_CUDA.
 */
#ifndef AFLAK_B200_H
#define AFLAK_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define AFK_ABI_VERSION 1

/* ---- status: IOErr variants (lib.rs:109-116) + two conditions the reference expresses as panics/CUDA ---- */
typedef enum afk_status {
    AFK_OK = 0,
    AFK_ERR_IO = 1,               /* IOErr::IoError */
    AFK_ERR_RAWLOADER = 2,        /* IOErr::RawLoaderError */
    AFK_ERR_FITS = 3,             /* IOErr::FITSErr */
    AFK_ERR_UNEXPECTED_INPUT = 4, /* IOErr::UnexpectedInput(String); message wording as precond.rs */
    AFK_ERR_SHAPE = 5,            /* IOErr::ShapeError */
    AFK_ERR_CUDA = 6,             /* surfaced to cake as UnexpectedInput("CUDA: ...") */
    AFK_ERR_PANIC = 7,            /* the reference would panic here (cake_fn! "Unexpected argument!", OOB index) */
    AFK_ERR_ARGUMENT = 8          /* cake ArgumentError::ConversionError (transform.rs:348-361) */
} afk_status;

/* Message of the last failing call on this thread (valid until the next call on the same thread). */
const char *afk_last_error(void);
int afk_abi_version(void);
/* number of CUDA devices visible; 0 means every compute entry point will fail with AFK_ERR_CUDA */
int afk_device_count(void);
int afk_set_device(int device);
/* Use an external stream (e.g. the caller's torch stream) for this thread on the current device; NULL restores
 * the thread's own stream. */
int afk_set_stream(void *cuda_stream);
void *afk_get_stream(void);
/* waits for this thread's stream; AFK_ERR_PANIC when a display chain enqueued by afk_colormap hit the reference's
 * tone-table panic since the last synchronisation */
int afk_synchronize(void);
/* kernels launched by this library since load (all threads) */
uint64_t afk_launch_count(void);

/* ---- type ids: the IOValue variant names (variant_name/src/lib.rs:24-29) ---- */
typedef enum afk_type {
    AFK_T_INTEGER = 0,
    AFK_T_FLOAT = 1,
    AFK_T_FLOAT2 = 2,
    AFK_T_FLOAT3 = 3,
    AFK_T_FLOAT3X3 = 4,
    AFK_T_STR = 5,
    AFK_T_BOOL = 6,
    AFK_T_FITS = 7,
    AFK_T_IMAGE = 8,
    AFK_T_MAP2DTO3DCOORDS = 9,
    AFK_T_ROI = 10,
    AFK_T_PATHS = 11,
    AFK_T_TONECURVE = 12,
    AFK_T_COUNT = 13
} afk_type;
const char *afk_type_name(afk_type t);          /* "Integer", "Float", ... */
int afk_type_from_name(const char *name);       /* -1 when unknown */
/* ConvertibleVariants::convertible (aflak_cake/src/lib.rs:100-112; table lib.rs:795-823) */
int afk_type_convertible(afk_type from, afk_type into);
/* EditableVariants (lib.rs:778-793) */
int afk_type_editable(afk_type t);

/* ---- images: device-resident WcsArray (unit.rs:203-208) ---- */
typedef struct afk_image afk_image;
typedef struct afk_fits afk_fits;
#define AFK_MAX_NDIM 4

/* Upload a host array (C order, shape[0] slowest: [NAXISn..NAXIS1], fits.rs:42-49) to the current device.
 * The copy is asynchronous when `host` is pinned; the returned handle has refcount 1. */
afk_image *afk_image_upload(const float *host, int ndim, const int64_t *shape);
/* Uninitialised device image. */
afk_image *afk_image_alloc(int ndim, const int64_t *shape);
/* Wrap device memory owned by the caller (e.g. a torch tensor); never freed by the library. */
afk_image *afk_image_wrap_device(float *device_ptr, int ndim, const int64_t *shape);
/* Fill with the synthetic cube generator (DESIGN.md "Synthetic inputs"): rows [y0,y0+rows) of a [nz,ny,nx] cube. */
afk_image *afk_image_synthetic(uint64_t seed, int64_t nz, int64_t ny, int64_t nx, int64_t y0, int64_t rows, int pathology);
void afk_image_retain(afk_image *img);   /* Clone of an Arc: O(1) */
void afk_image_release(afk_image *img);  /* last release frees stream-ordered */
int afk_image_ndim(const afk_image *img);
const int64_t *afk_image_shape(const afk_image *img);
int64_t afk_image_len(const afk_image *img);
float *afk_image_device_ptr(const afk_image *img);
int afk_image_device(const afk_image *img);
/* Blocking read-back (UI tooltips, FITS export): waits for the producer, copies to host. */
int afk_image_download(const afk_image *img, float *host);
/* metadata carried on the host side (unit.rs:190-242) */
const char *afk_image_unit(const afk_image *img);       /* "" == Unit::None */
int afk_image_set_unit(afk_image *img, const char *unit);
int afk_image_homogeneous(const afk_image *img);
const char *afk_image_tag(const afk_image *img);        /* NULL == visualization: None */
int afk_image_set_tag(afk_image *img, const char *tag);
int afk_image_has_meta(const afk_image *img);
/* axis i of MetaWcsArray.axes (FITS order: axis 0 = NAXIS1): name/unit, "" when absent */
int afk_image_set_axis(afk_image *img, int axis, const char *name, const char *unit);
/* linear world coordinates of FITS axis i (MetaWcsArray.wcs, unit.rs:210-214): world = CRVAL + CDELT * (pixel + 1 - CRPIX),
 * evaluated in f64 and returned as f32; pixel is 0-based.  Restates fitrs 0.5.0 `WCS` for axis-aligned headers. */
int afk_image_set_wcs_axis(afk_image *img, int axis, double crval, double crpix, double cdelt);
/* WcsArray::pix2world (unit.rs:335-342): returns 1 and writes *world, or 0 when the image carries no metadata (None) */
int afk_image_pix2world(const afk_image *img, int axis, float pixel, float *world);
const char *afk_image_axis_name(const afk_image *img, int axis);
const char *afk_image_axis_unit(const afk_image *img, int axis);

/* ---- FITS ingest: WcsArray::from_hdu (unit.rs:283-331) + FitsDataToArray (fits.rs:42-103) over fitrs 0.5.0.
 *      The header is parsed on the host; the data unit crosses PCIe as the raw big-endian bytes of the file and is
 *      decoded on the device (byte swap, i16/i32/u32/f64 -> f32, BLANK -> NaN, then value * BSCALE + BZERO in two
 *      rounded f32 steps), chunk by chunk, each decode overlapping the next chunk's copy. ---- */
typedef enum afk_fits_format {
    AFK_FITS_F32 = 0, /* BITPIX -32  FitsData::FloatingPoint32 */
    AFK_FITS_F64 = 1, /* BITPIX -64  FitsData::FloatingPoint64, `f as f32` */
    AFK_FITS_I16 = 2, /* BITPIX  16  FitsData::IntegersI32 */
    AFK_FITS_I32 = 3, /* BITPIX  32  FitsData::IntegersI32 */
    AFK_FITS_U32 = 4  /* BITPIX  32 read as unsigned: FitsData::IntegersU32 */
} afk_fits_format;
/* Decode a data unit that is already in host memory (pinned memory makes the copies asynchronous).  shape as
 * afk_image_upload: [NAXISn..NAXIS1].  blank is only read when has_blank (integer formats). */
afk_image *afk_image_ingest(const void *raw_be, afk_fits_format fmt, int ndim, const int64_t *shape, int has_blank,
                            int64_t blank, float bscale, float bzero);
/* Same for one spatial band of a [nz, ny, nx] cube (rows [y0, y0+rows) of every frame): the per-GPU upload of a
 * sharded cube (afk_shard_band).  The result has shape [nz, rows, nx]. */
afk_image *afk_image_ingest_band(const void *raw_be, afk_fits_format fmt, int64_t nz, int64_t ny, int64_t nx, int64_t y0,
                                 int64_t rows, int has_blank, int64_t blank, float bscale, float bzero);
/* fitrs::Fits::open (lib.rs:826-835): parse every HDU header; the data units stay on disk.  NULL on failure
 * (afk_last_error() has the reference's "Could not open file ..." wording). */
afk_fits *afk_fits_open(const char *path);
void afk_fits_retain(afk_fits *f);
void afk_fits_release(afk_fits *f);
size_t afk_fits_hdu_count(const afk_fits *f);
/* value of a header card as text (strings unquoted and right-trimmed), NULL when the key is absent */
const char *afk_fits_header(const afk_fits *f, size_t hdu, const char *key);
/* run_fits_to_image (lib.rs:883-910): HDU by EXTNAME, else by index; reads the data unit through a pinned double
 * buffer and decodes it on the device; attaches BUNIT, CTYPEn/CUNITn and the linear WCS (CRVALn/CRPIXn/CDELTn). */
afk_status afk_fits_to_image(const afk_fits *f, int64_t hdu_idx, const char *extension, afk_image **out);
/* same for this rank's band of a 3-D HDU (rows split with afk_shard_band): each GPU reads only its own rows */
afk_status afk_fits_to_image_band(const afk_fits *f, int64_t hdu_idx, const char *extension, int world, int rank, afk_image **out);
/* timings of the last ingest on this thread, in ms: [0] file/host read into pinned memory, [1] whole call (wall),
 * [2] device span from the first copy to the last decode (CUDA events); bytes = raw bytes moved over PCIe */
int afk_ingest_last_stats(double ms[3], uint64_t *bytes);

/* ---- values: IOValue (lib.rs:66-83), refcounted like the Arc<T> the cake cache stores ---- */
typedef struct afk_value afk_value;
afk_value *afk_value_integer(int64_t v);
afk_value *afk_value_float(float v);
afk_value *afk_value_float2(const float v[2]);
afk_value *afk_value_float3(const float v[3]);
afk_value *afk_value_float3x3(const float v[9]);
afk_value *afk_value_str(const char *s);
afk_value *afk_value_bool(int b);
afk_value *afk_value_image(afk_image *img); /* retains img */
afk_value *afk_value_fits(afk_fits *f);     /* retains f: IOValue::Fits(Arc<fitrs::Fits>) */
afk_fits *afk_value_get_fits(const afk_value *v); /* borrowed */
afk_value *afk_value_paths(const char *const *paths, size_t n); /* IOValue::Paths(PATHS::FileList) */
size_t afk_value_paths_count(const afk_value *v);
const char *afk_value_paths_at(const afk_value *v, size_t i);
/* roi::ROI (roi.rs:5-10): All, or PixelList of (i, j) = (x, y) pairs: xy holds 2 * n_pixels values */
afk_value *afk_value_roi_all(void);
afk_value *afk_value_roi_pixels(const uint64_t *xy, size_t n_pixels);
int afk_value_roi_is_all(const afk_value *v);
size_t afk_value_roi_count(const afk_value *v);
const uint64_t *afk_value_roi_data(const afk_value *v);
/* Map2dTo3dCoords: Array2<[f32; 3]> (lib.rs:79), [rows][cols][3] */
afk_value *afk_value_map2dto3d(const float *xyz, int64_t rows, int64_t cols);
const float *afk_value_get_map(const afk_value *v, int64_t *rows, int64_t *cols);
/* ToneCurveState: only `arr` is read by operators (imgui_tone_curve/src/lib.rs:10-48); cp = n_cp (x,y) pairs */
afk_value *afk_value_tone_curve(const float *arr, size_t n, const float *cp_xy, size_t n_cp);
/* UiToneCurve::create_curve (imgui_tone_curve/src/lib.rs:68-158): the table `arr` the widget builds from n_cp control
 * points (x,y pairs).  *n_out = number of entries (256 for 2 or 3 points, a spline-dependent count otherwise, 0 for fewer
 * than 2); arr_out may be NULL to query the size.  A NaN abscissa is the reference's unwrap() panic (AFK_ERR_PANIC). */
afk_status afk_tone_curve_from_points(const float *cp_xy, size_t n_cp, float *arr_out, size_t cap, size_t *n_out);
/* DefaultFor::default_for (lib.rs:760-776); NULL for types without a default (Image, Fits, Map2dTo3dCoords) */
afk_value *afk_value_default_for(afk_type t);
void afk_value_retain(afk_value *v);
void afk_value_release(afk_value *v);
afk_type afk_value_type(const afk_value *v);
int afk_value_get_integer(const afk_value *v, int64_t *out);
int afk_value_get_float(const afk_value *v, float *out);
int afk_value_get_floats(const afk_value *v, float *out, size_t cap); /* Float2/3/3x3/ToneCurve arr; returns count */
int afk_value_get_bool(const afk_value *v, int *out);
const char *afk_value_get_str(const afk_value *v);
afk_image *afk_value_get_image(const afk_value *v); /* borrowed */
/* ConvertibleVariants::convert (aflak_cake/src/lib.rs:86-98): Integer->Float (`as f32`), Float->Integer (round) */
afk_value *afk_value_convert(const afk_value *v, afk_type into);
/* PartialEq for IOValue (lib.rs:85-104); images compare shape, unit, tag and device contents */
int afk_value_eq(const afk_value *a, const afk_value *b);

/* ---- registry: TRANSFORMATIONS (lib.rs:141-747) ---- */
typedef struct afk_result {
    afk_status status;  /* Result<IOValue, IOErr> discriminant */
    afk_value *value;   /* Ok(v): refcount 1, owned by the caller */
    char *message;      /* Err(e): e's payload string (malloc'ed; free with afk_result_clear) */
} afk_result;
void afk_result_clear(afk_result *r);
/* Display for IOErr (lib.rs:118-130): "Unexpected input! ..." etc.  Returns a malloc'ed string. */
char *afk_result_display(const afk_result *r);
void afk_free(void *p);

/* PlainFunction = fn(Vec<Bow<T>>) -> Vec<Result<T,E>> (transform.rs:78): one result per declared output. */
typedef afk_status (*afk_plain_function)(const afk_value *const *inputs, size_t n_inputs, afk_result *outputs, size_t n_outputs);

typedef struct afk_input_slot { /* TransformInputSlot (transform.rs:64-70) */
    const char *name;
    afk_type type_id;
    const afk_value *default_value; /* NULL when the slot has no default */
} afk_input_slot;

typedef struct afk_transform_desc { /* Algorithm::Function (transform.rs:33-48) */
    const char *name;        /* FnTransformId */
    const char *kind;
    const char *description;
    uint8_t major, minor, patch;
    size_t n_inputs;
    const afk_input_slot *inputs;
    size_t n_outputs;
    const afk_type *outputs;
    afk_plain_function f;
} afk_transform_desc;

size_t afk_transform_count(void);
const afk_transform_desc *afk_transform_at(size_t i);
/* NamedAlgorithms::get_transform (lib.rs:749-758) */
const afk_transform_desc *afk_get_transform(const char *name);
/* TransformCaller feed + call (transform.rs:324-397): type-check each argument against the slot, convert
 * Integer<->Float where the table allows, then call f.  NULL inputs[i] takes the slot default. */
afk_status afk_transform_call(const afk_transform_desc *t, const afk_value *const *inputs, size_t n_inputs,
                              afk_result *outputs, size_t n_outputs);

/* Reduction arithmetic: 0 = compensated (default; <=1e-5 of the f64 sum), 1 = sequential (bit-exact with the
 * reference's left-to-right f32 sum; slower on small images because it cannot split the spectral axis).
 * cake calls operators from arbitrary rayon workers (aflak_cake/src/dst/compute.rs:206-224), so the mode is a
 * process-wide default (afk_set_default_sum_mode, set once by the host) with an optional per-thread override
 * (afk_set_sum_mode; -1 removes the override).  afk_get_sum_mode returns the mode in force on the calling thread. */
int afk_set_default_sum_mode(int mode);
int afk_set_sum_mode(int mode);
int afk_get_sum_mode(void);
/* Return the cached (freed) blocks of the current device's stream-ordered pools to the driver; synchronises the device.
 * For hosts about to make an allocation that needs nearly the whole device (the 128 GiB cube of BASELINE configs[4]). */
int afk_trim_pools(void);

/* ---- fused paths the evaluator substitutes for recognised sub-graphs (unfused operators remain the parity
 *      reference).  Inputs are borrowed, outputs have refcount 1. ---- */
/* average(cube,s0,e0) / average(cube,s1,e1)  -- BASELINE config 3 */
afk_status afk_fused_average_ratio(const afk_image *cube, int64_t s0, int64_t e0, int64_t s1, int64_t e1, afk_image **out);
/* convert_to_logscale(img, a, min, max) with (min, max) = image_min_max(img) (lib.rs:180-197, 282-290, 1598-1611; the tail of
 * the equivalent-width template): two kernels, the two floats stay on the device */
afk_status afk_fused_logscale_minmax(const afk_image *img, float a, afk_image **out);
/* equivalent-width template (src/templates.rs:143-463): three averages -> ratio_from_bands ->
 * linear_composition -> create_equivalent_width; minmax (may be NULL) receives image_min_max of the result */
afk_status afk_fused_equivalent_width(const afk_image *cube, int64_t on_s, int64_t on_e, int64_t off1_s, int64_t off1_e,
                                      int64_t off2_s, int64_t off2_e, float fl, float max, int is_emission,
                                      afk_image **out, float minmax[2]);

/* ---- display path: ColorLUT (lut.rs) + make_raw_image (image.rs:20-84) ---- */
typedef struct afk_lut afk_lut;
afk_lut *afk_lut_linear(const float *stop_values, const uint8_t *stop_rgb, size_t n_stops); /* ColorLUT::linear */
afk_lut *afk_lut_builtin(const char *name); /* BuiltinLUT::name(): "Grey", "Flame", "HeatMap_Inv", ... */
size_t afk_lut_builtin_count(void);
const char *afk_lut_builtin_name(size_t i);
void afk_lut_release(afk_lut *lut);
void afk_lut_set_lims(afk_lut *lut, float min, float mid, float max); /* set_lims (lut.rs:232-236) */
void afk_lut_lims(const afk_lut *lut, float out[3]);
int afk_lut_set_gradient(afk_lut *lut, const float *stop_values, const uint8_t *stop_rgb, size_t n_stops);
void afk_lut_color_at(const afk_lut *lut, float point, uint8_t rgb[3]);                       /* lut.rs:137-149 */
void afk_lut_color_at_bounds(const afk_lut *lut, float point, float vmin, float vmax, uint8_t rgb[3]);
/* copy of the 65536x3 table (host) */
void afk_lut_table(const afk_lut *lut, uint8_t *out_rgb);
/* bounds() iterator (lut.rs:196-200): writes up to cap pairs, returns the number of pairs */
size_t afk_lut_bounds(const afk_lut *lut, float *v1, uint8_t *c1, float *v2, uint8_t *c2, size_t cap);

/* The display kernels (apply_tone_curve, colour map, fused chain) replace divisions by launch-uniform divisors with an
 * exact three-instruction form and evaluate the chain's tone index with a guard band; both give the
 * reference's indices and bytes.  on = 1 makes every pixel take the reference's own IEEE sequence instead (also
 * AFK_DISPLAY_EXACT=1 in the environment): the baseline the parity tests compare the fast forms with. */
int afk_set_display_exact(int on);
int afk_get_display_exact(void);

/* lims::get_vmin / get_vmax (aflak_plot/src/lims.rs:7-35): NaN-skipping, NaN when every pixel is NaN */
afk_status afk_image_vmin_vmax(const afk_image *img, float *vmin, float *vmax);

/* ---- display statistics `Image::new` computes beside the colour mapping (aflak_plot/src/imshow/image.rs:155-176) ---- */
/* hist::histogram (hist.rs:10-36): 100 bins over [vmin, vmax]; every pixel is counted (NaN and values below vmin in
 * bin 0, values at or above vmax in bin 99).  starts / ends (may be NULL) receive Bin.start / Bin.end. */
afk_status afk_image_histogram(const afk_image *img, float vmin, float vmax, uint64_t counts[100], float starts[100], float ends[100]);
/* hist::histogram_color (hist.rs:38-66): img is [3, ny, nx]; 256 bins; counts[bin * 3 + channel] */
afk_status afk_image_histogram_color(const afk_image *img, float vmin, float vmax, uint64_t counts[768], float starts[256], float ends[256]);
/* lims::get_vmed_normalized and get_vmad_normalized(image, vmed) (lims.rs:37-106): median of (v - vmin) / (vmax - vmin)
 * over the pixels where that is finite, and the median absolute deviation from it.  Exact order statistics.
 * Errors (AFK_ERR_UNEXPECTED_INPUT): "Empty data!, med", "Empty data!, mad". */
afk_status afk_image_vmed_vmad(const afk_image *img, float *vmed, float *vmad);
/* Image::zscale(nsamples, contrast) (zscale.rs:45-100; the UI calls it with (1000, 0.25)): *has_limits = 0 is the
 * reference's (None, None) */
afk_status afk_image_zscale(const afk_image *img, int nsamples, float contrast, int *has_limits, float *z1, float *z2);

typedef struct afk_colormap_chain { /* optional fused upstream chain (BASELINE config 4) */
    float logscale_a, logscale_vmin, logscale_vmax; /* convert_to_logscale; vmin = vmax = NaN: use image_min_max of the image
                                                     * (lib.rs:180-197), computed on the device with no read-back in between */
    float normalize_min, normalize_max;             /* normalize_image */
    const float *tone_table;                        /* apply_tone_curve (host pointer) */
    size_t tone_len;
} afk_colormap_chain;

/* make_raw_image on the vertically flipped image.  vmin/vmax NaN -> computed with afk_image_vmin_vmax.
 * channels: 3 = packed RGB8 (the reference's ClientFormat::U8U8U8), 4 = RGBA8 (A=255).
 * out_device: device buffer of ny*nx*channels bytes (headless stand-in for the GL texture).
 * Stream-ordered, no host synchronisation.  With a chain, a pixel whose tone-curve index lies beyond the table (the
 * reference's apply_tone_curve panics there, lib.rs:1866) is mapped with the last table entry and the panic is reported
 * as AFK_ERR_PANIC by the next afk_synchronize() on this thread (afk_colormap_host reports it itself). */
afk_status afk_colormap(const afk_image *img, const afk_lut *lut, float vmin, float vmax, int flip_y, int channels,
                        const afk_colormap_chain *chain, uint8_t *out_device);
/* same, result copied to host memory */
afk_status afk_colormap_host(const afk_image *img, const afk_lut *lut, float vmin, float vmax, int flip_y, int channels,
                             const afk_colormap_chain *chain, uint8_t *out_host);
/* make_raw_image_rgb (image.rs:51-84): img is [3,ny,nx], one LUT and one (vmin,vmax) per channel */
afk_status afk_colormap_rgb_host(const afk_image *img, const afk_lut *const lut[3], const float vmin[3], const float vmax[3],
                                 int flip_y, int channels, uint8_t *out_host);

/* ---- raw kernel layer: device pointers + the thread's stream, no handles (bench.py, multi-GPU bands) ---- */
/* epilogue: 0 sum, 1 mean, 2 mean0/mean1, 3 EW.  params is read for EW only: exactly 5 floats (a, b, fl, max, sign).
 * `out` may be a peer-mapped pointer. */
afk_status afk_k_reduce(const float *cube, int64_t nz, int64_t n2, int nwin, const int64_t *starts, const int64_t *ends,
                        int epilogue, const float *params, int sum_mode, float *out);
/* afk_k_reduce for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 * No counterpart in the reference.  Same arguments, arithmetic, kernels and
 * bits as afk_k_reduce, but the split-K scratch (f64 partials + arrival counters) lives in a file-scope __device__ array of
 * 32.25 MiB that every GPU holds from library load to exit; the call performs no cudaMalloc / cudaFree / pool allocation.
 * cuda_stream: the cudaStream_t to launch on (NULL = the legacy default stream); `cube` and `out` must be ready on it.
 * Limits: a launch plan that needs more than 32 MiB of partials or more than 65536 arrival counters (e.g. the three-window
 * EW epilogue on images of a megapixel and more) is rejected with AFK_ERR_UNEXPECTED_INPUT -- use afk_k_reduce there.
 * All calls on one device share the array: calls that overlap on the GPU race (serialise them); the counters are cleared
 * on the stream in every call.  The first call on a device resolves the array's address: make it outside a capture. */
afk_status afk_k_reduce_static(const float *cube, int64_t nz, int64_t n2, int nwin, const int64_t *starts, const int64_t *ends,
                               int epilogue, const float *params, int sum_mode, float *out, void *cuda_stream);
/* text description of the kernel instantiation and launch shape afk_k_reduce uses for this problem (reports) */
int afk_k_reduce_describe(const float *cube, int64_t n2, int nwin, const int64_t *starts, const int64_t *ends, int sum_mode,
                          const float *out, char *buf, size_t cap);
/* the ingest decode kernel alone: device raw bytes -> device f32 (out may alias raw for 4-byte formats) */
afk_status afk_k_ingest(const void *device_raw_be, afk_fits_format fmt, int64_t n, int has_blank, int64_t blank, float bscale,
                        float bzero, float *device_out);
/* CUDA-event timing of reduce-kernel launches on this thread: enable, run, then read (sum of ms, launches). */
int afk_profile_reduce(int enable);
int afk_profile_reduce_read(double *total_ms, uint64_t *launches);

/* CUDA IPC helpers for the fused gather (one process per GPU): export this image's allocation, open a peer's. */
int afk_ipc_export(const afk_image *img, uint8_t handle[64]);
float *afk_ipc_open(const uint8_t handle[64]);
int afk_ipc_close(float *peer_ptr);

/* ---- spatial sharding across the GPUs of one box + fused gather of the 2-D result (one process per GPU).
 *      The reference has no distributed path (single process, rayon: aflak_cake/src/dst/compute.rs:152-224); this is
 *      the B200 form of "large cubes shard by spatial tiles, the only NVLink traffic is the gather of the 2-D result".
 *      Protocol: every rank calls afk_gather_create, the 64-byte handles are exchanged by the host's process group
 *      (world*64 bytes, rank-major), every rank calls afk_gather_connect; then per step afk_gather_reduce (the reduce
 *      kernel stores its band into every rank's full image through peer mappings and publishes an epoch flag) and
 *      afk_gather_wait (stream-ordered one-warp wait for all bands).  Results alternate between two slots, so step
 *      n+1 never overwrites an image a peer may still be reading from step n.  The library enforces what that relies
 *      on: every afk_gather_reduce must be followed by an afk_gather_wait before the next one, both on the stream the
 *      gather was created on (anything else fails with AFK_ERR_UNEXPECTED_INPUT); operators that read the handle of
 *      afk_gather_result_image from other streams are recorded and the next reduce is ordered after them.  A band that
 *      does not arrive within the time-out is overwritten with NaN and reported by afk_gather_check. ---- */
typedef struct afk_gather afk_gather;
/* y-band [y0, y0+rows) of `rank`: rows split evenly, the first ny % world ranks get one extra row */
int afk_shard_band(int64_t ny, int world, int rank, int64_t *y0, int64_t *rows);
afk_gather *afk_gather_create(int world, int rank, int64_t ny_total, int64_t nx, uint8_t handle_out[64]);
int afk_gather_connect(afk_gather *g, const uint8_t *handles);
/* The same gather over memory the HOST supplies (symmetric allocation: torch symmetric memory, NVSHMEM, or cuMem + a multicast
 * object): afk_gather_bytes() zeroed bytes on every rank; peer_bases[world] = every rank's buffer as mapped in this process
 * (peer_bases[rank] is the local one); multicast_base = NVLink-SHARP multicast mapping of the same buffers, or NULL.  With a
 * multicast mapping the reduce kernel's epilogue and the RGBA band broadcast emit ONE multimem.st per 16 bytes instead of one
 * store per GPU.  No afk_gather_connect needed; afk_gather_destroy does not free the memory. */
size_t afk_gather_bytes(int64_t ny_total, int64_t nx);
afk_gather *afk_gather_create_symmetric(int world, int rank, int64_t ny_total, int64_t nx, void *const *peer_bases, void *multicast_base);
int afk_gather_band(const afk_gather *g, int rank, int64_t *y0, int64_t *rows);
/* band_cube: device pointer to this rank's [nz, rows, nx] band; epilogue/sum_mode as afk_k_reduce; params is read for
 * epilogue 3 (EW) only and then holds exactly 5 floats (a, b, fl, max, sign) */
afk_status afk_gather_reduce(afk_gather *g, const float *band_cube, int64_t nz, int nwin, const int64_t *starts,
                             const int64_t *ends, int epilogue, const float *params, int sum_mode);
/* *result = device pointer to this rank's copy of the full [ny_total, nx] image, valid once the stream reaches it */
afk_status afk_gather_wait(afk_gather *g, float **result, double timeout_s);
/* host-synchronising: AFK_ERR_CUDA if a wait timed out (a peer's band never arrived) */
afk_status afk_gather_check(afk_gather *g);
/* the current full image as an Image handle that borrows the slot's memory (release it like any handle; the pixels are
 * valid until the second afk_gather_reduce from now), for further operators (LUT, min/max, ...).  Call after afk_gather_wait. */
afk_image *afk_gather_result_image(afk_gather *g);
/* Sharded display path.  (1) Min/max exchange: every afk_gather_reduce also finds the range of its band and
 * sends the two floats with its flag; afk_gather_wait combines them, so the handle returned by afk_gather_result_image
 * carries the whole image's range and image_min_max (aflak_primitives/src/lib.rs:180-197), convert_to_logscale fed by it
 * (:1598-1611), get_vmin / get_vmax (aflak_plot/src/lims.rs:7-35) and the colour mapping make no pass over the gathered
 * image.  (2) RGBA band gather: afk_gather_colormap maps THIS RANK'S BAND of the current gathered image (make_raw_image,
 * aflak_plot/src/imshow/image.rs:20-49; limits from the whole image when vmin/vmax or the chain's logscale range are NaN)
 * and stores the RGBA8 band into every rank's full [ny, nx, 4] byte image; afk_gather_wait_rgba enqueues the wait and
 * returns this rank's copy.  Call order per step: reduce, wait, colormap, wait_rgba, all on the gather's stream.  nx must
 * be a multiple of 4. */
afk_status afk_gather_colormap(afk_gather *g, const afk_lut *lut, float vmin, float vmax, int flip_y, const afk_colormap_chain *chain);
afk_status afk_gather_wait_rgba(afk_gather *g, uint8_t **rgba, double timeout_s);
void afk_gather_destroy(afk_gather *g);

/* ---- mini evaluator with a device-handle cache: DST::compute + Cache (dst/compute.rs:153-274, cache.rs) ---- */
typedef struct afk_graph afk_graph;
afk_graph *afk_graph_new(void);
void afk_graph_free(afk_graph *g);
/* add_transform / add_constant: returns the node index (t_idx) or -1 */
int64_t afk_graph_add_transform(afk_graph *g, const char *name);
int64_t afk_graph_add_constant(afk_graph *g, const afk_value *v);
/* replace a constant's value (node_editor::update_constant_node): bumps updated_on */
int afk_graph_set_constant(afk_graph *g, int64_t t_idx, const afk_value *v);
/* connect(Output(t_idx,out_i) -> Input(t_idx,in_i)); type-checked (dst/build.rs:45-88) */
afk_status afk_graph_connect(afk_graph *g, int64_t from_t, size_t out_i, int64_t to_t, size_t in_i);
afk_status afk_graph_disconnect(afk_graph *g, int64_t to_t, size_t in_i);
/* compute Output(t_idx,out_i) with memoisation: a node is recomputed iff its cached time < updated_on of the
 * node or any ancestor (cache.rs:130-136, dst/mod.rs:65-73).  Result value refcount +1 for the caller. */
afk_status afk_graph_compute(afk_graph *g, int64_t t_idx, size_t out_i, afk_result *out);
/* number of operator calls actually executed (cache misses) since creation */
uint64_t afk_graph_calls(const afk_graph *g);
/* enable/disable substitution of fused kernels for recognised sub-graphs (default on) */
void afk_graph_set_fusion(afk_graph *g, int enable);

#ifdef __cplusplus
}
#endif
#endif /* AFLAK_B200_H */
