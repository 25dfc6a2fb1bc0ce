/*
 * aflak_oracle.h -- CPU restatement of aflak's data-parallel hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product:
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library, and only as the checker or the
 * timed CPU baseline.  The product (aflak_b200/) never links or calls it.
 *
 * Parity status: the reference (pure Rust, /root/reference/src) cannot be
 * compiled here (no rustc/cargo, crates not vendored).  This file restates the
 * algorithm step by step in strict f32 (build with -ffp-contract=off, no
 * -ffast-math).  Pinned parts: the colour LUT (11 golden RGB triples and 4
 * bounds() tuples of aflak_plot/src/imshow/lut.rs:292-343).  Reductions and
 * elementwise operators have no value-asserting reference test and their inner
 * loops live in ndarray 0.12.1 (Cargo.lock:1112, not on disk):
 * "parity unpinned" for those -- anchored on hand-derivable known answers.
 *
 * All file:line citations are relative to /root/reference/src/.
 */
#ifndef AFLAK_ORACLE_H
#define AFLAK_ORACLE_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* return codes */
#define ORC_OK 0
#define ORC_ERR_UNEXPECTED_INPUT 1 /* IOErr::UnexpectedInput, message in orc_last_error() */
#define ORC_ERR_PANIC 2            /* the reference would panic (index out of bounds) */

const char *orc_last_error(void);

/* ---- axis-0 reductions (aflak_primitives/src/lib.rs:1303-1330) ---- */
int orc_integral(const float *cube, int64_t nz, int64_t n2, int64_t start, int64_t end, float *out);
int orc_average(const float *cube, int64_t nz, int64_t n2, int64_t start, int64_t end, float *out);
/* [nz,1]-shaped input: ndarray 0.12.1 takes the stride-1 branch (8-way unrolled fold) */
int orc_integral_lane(const float *col, int64_t nz, int64_t start, int64_t end, float *out);
/* f64 oracles used for the tolerance test: sum and sum of |x| per pixel */
void orc_sum_f64(const float *cube, int64_t n2, int64_t start, int64_t end, double *sum, double *abs_sum);
/* second/third outputs of integral/average (lib.rs:393-395) */
void orc_middle_width(int64_t start, int64_t end, float *middle, float *width);
/* all-cores variants for the timed CPU baseline (same arithmetic per pixel) */
int orc_integral_mt(const float *cube, int64_t nz, int64_t n2, int64_t start, int64_t end, float *out, int threads);
int orc_average_mt(const float *cube, int64_t nz, int64_t n2, int64_t start, int64_t end, float *out, int threads);

/* ---- FITS data unit -> f32 (unit.rs:283-331, fits.rs:42-103); fmt: 0 f32, 1 f64, 2 i16, 3 i32, 4 u32 ---- */
int orc_fits_decode(const uint8_t *raw, int fmt, int64_t n, int has_blank, int64_t blank, float bscale, float bzero, float *out);

/* ---- other axis-0 reductions (lib.rs:1332-1582) ---- */
int orc_variance(const float *cube, int64_t nz, int64_t n2, int64_t start, int64_t end, int take_sqrt, float *out);
void orc_variance_f64(const float *cube, int64_t n2, int64_t start, int64_t end, double *out);
int orc_median(const float *cube, int64_t nz, int64_t n2, int64_t start, int64_t end, float *out);
int orc_minmax_fold(const float *cube, int64_t nz, int64_t n2, int64_t start, int64_t end, int is_min, float *out);
/* wcs = {CRVAL, CRPIX, CDELT} of the axis, or NULL for "pixel maps to itself" */
float orc_pix2world(const double *wcs, float pixel);
int orc_argmap(const float *cube, int64_t nz, int64_t n2, int64_t start, int64_t end, int64_t range, int is_min,
               const double *wcs, float *out);
int orc_centroid(const float *cube, int64_t nz, int64_t n2, int64_t start, int64_t end, const double *wcs, float *out);
void orc_centroid_f64(const float *cube, int64_t n2, int64_t start, int64_t end, const double *wcs, double *out);
int orc_centroid_mask(const float *cube, int64_t nz, int64_t n2, const float *start_mask, const float *end_mask,
                      const double *wcs, float *out);

/* ---- slicing / gather (lib.rs:912-925, 1894-1962) ---- */
int orc_slice_one_frame(const float *cube, int64_t nz, int64_t n2, int64_t frame, float *out);
int orc_down_sampling(const float *img, int ndim, const int64_t *shape, int64_t n, float *out, int64_t *out_shape);

/* ---- elementwise (lib.rs:1288-1297, 1584-1638, 1858-1871, 180-197, 226-246) ---- */
void orc_linear_composition(const float *u, const float *v, float a, float b, int64_t n, float *out);
void orc_ratio(const float *u, const float *v, int64_t n, float *out);
void orc_image_multiplier(const float *u, const float *v, float a, float b, int64_t n, float *out);
void orc_equivalent_width(const float *off, const float *on, float fl, float max, int is_emission, int64_t n, float *out);
void orc_image_min_max(const float *img, int64_t n, float *vmin, float *vmax);
void orc_convert_to_logscale(const float *img, int64_t n, float a, float vmin, float vmax, float *out);
void orc_log10(const float *img, int64_t n, float *out);
int orc_apply_tone_curve(const float *img, int64_t n, const float *table, int64_t table_len, float *out);
int orc_ratio_from_bands(float z, float z1, float z2, float *off_ratio_2, float *off_ratio_1);
/* all-cores elementwise (the reference uses rayon par_iter_mut here) */
void orc_convert_to_logscale_mt(const float *img, int64_t n, float a, float vmin, float vmax, float *out, int threads);

/* ---- tone-curve table (imgui_tone_curve/src/lib.rs:21-48,68-158) ---- */
/* returns table length (<= cap) or -1 when cap is too small */
int64_t orc_tone_curve_default(float *table, int64_t cap);
int64_t orc_create_curve(const float *cp_xy, int64_t n_cp, float *table, int64_t cap);

/* ---- colour LUT (aflak_plot/src/imshow/lut.rs, image.rs, lims.rs, util.rs) ---- */
#define ORC_LUT_SIZE 65536
/* stops: n entries of (value, r, g, b); lims = (min, mid, max) in force at build time */
void orc_lut_build(const float *stop_v, const uint8_t *stop_rgb, int64_t n_stops, const float lims[3], uint8_t *lut /* 65536*3 */);
/* bounds(): writes up to cap pairs ((v1,c1),(v2,c2)); returns number of pairs */
int64_t orc_lut_bounds(const float *stop_v, const uint8_t *stop_rgb, int64_t n_stops, const float lims[3],
                       float *v1, uint8_t *c1, float *v2, uint8_t *c2, int64_t cap);
float orc_mtf(float m, float x);
uint32_t orc_lut_index(float point, const float lims[3]);
void orc_color_at(const uint8_t *lut, const float lims[3], float point, uint8_t rgb[3]);
void orc_color_at_bounds(const uint8_t *lut, const float lims[3], float point, float vmin, float vmax, uint8_t rgb[3]);
/* lims.rs:7-35,108-124: returns 0 and writes value, 1 when the image is empty */
int orc_get_vmin(const float *img, int64_t n, float *vmin);
int orc_get_vmax(const float *img, int64_t n, float *vmax);
/* image.rs:20-49 on the vertically flipped view (image.rs:123-129); channels = 3 (RGB8) or 4 (RGBA8, A=255) */
int orc_make_raw_image(const float *img, int64_t ny, int64_t nx, float vmin, float vmax, const uint8_t *lut,
                       const float lims[3], int channels, uint8_t *out);
/* image.rs:51-84: img is [3,ny,nx]; lut3 = 3 tables, lims3 = 3x3; flipped like coerce_to_array_view3 */
int orc_make_raw_image_rgb(const float *img, int64_t ny, int64_t nx, const float vmin[3], const float vmax[3],
                           const uint8_t *lut3, const float *lims3, int channels, uint8_t *out);
int orc_make_raw_image_mt(const float *img, int64_t ny, int64_t nx, float vmin, float vmax, const uint8_t *lut,
                          const float lims[3], int channels, uint8_t *out, int threads);

/* ---- spatial reductions / gathers and remaining elementwise operators (lib.rs:509-631, 928-1169, 1233-1256, 1795-1892) ---- */
int orc_extract_wave(const float *cube, int64_t nz, int64_t ny, int64_t nx, int roi_all, const uint64_t *xy, int64_t npix,
                     float *out, double *out64);
int orc_extrude(const float *cube, int64_t nz, int64_t ny, int64_t nx, int roi_all, const uint64_t *xy, int64_t npix, float *out);
void orc_make_plane3d(const float p0[3], const float d1[3], const float d2[3], int64_t count1, int64_t count2, float *map);
int orc_slice_3d_to_2d(const float *cube, int64_t d0, int64_t d1, int64_t d2, const float *map, int64_t n, float *out);
void orc_velocity_field(const float *img, int64_t n, float w0, float *out);
void orc_histogram_transformation(const float *img, int64_t n, const float parameters[3], float *out);
void orc_arcsinh_stretch(const float *img, int64_t n, float beta, float *out);
void orc_color_image_to_hsv(const float *img, int64_t n, float *out);

/* ---- display statistics (aflak_plot/src/imshow/hist.rs, lims.rs:37-106, zscale.rs) ---- */
void orc_histogram(const float *img, int64_t n_per_channel, int channels, float min, float max, int nb, uint64_t *counts);
void orc_histogram_bins(float min, float max, int nb, float *starts, float *ends);
int orc_vmed_normalized(const float *img, int64_t n, float vmin, float vmax, float *out);
int orc_vmad_normalized(const float *img, int64_t n, float vmin, float vmax, float vmed, float *out);
int orc_zscale(const float *img, int64_t nc, int64_t nl, int nsamples, float contrast, int *has, float *z1, float *z2);

/* ---- synthetic cube generator shared with the device (DESIGN.md "Synthetic inputs") ---- */
/* fills out[nz, rows, nx] for rows [y0, y0+rows) of a cube whose full size is [nz, ny, nx] */
void orc_gen_cube(uint64_t seed, int64_t nz, int64_t ny, int64_t nx, int64_t y0, int64_t rows, int pathology,
                  float *out, int threads);
float orc_gen_voxel(uint64_t seed, int64_t nz, int64_t ny, int64_t nx, int64_t k, int64_t y, int64_t x, int pathology);

int orc_max_threads(void);

#ifdef __cplusplus
}
#endif
#endif
