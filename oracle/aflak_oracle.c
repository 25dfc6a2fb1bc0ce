/*
 * aflak_oracle.c -- CPU restatement of aflak's hot path (see aflak_oracle.h).
 *
 * TEST INFRASTRUCTURE ONLY; never linked into or called by the product.
 * Build: gcc -O2 -ffp-contract=off -fno-fast-math -fopenmp -shared -fPIC (oracle/Makefile).
 * Strict f32: every intermediate the reference rounds to f32 is rounded here
 * (no FMA contraction, no double temporaries unless the reference has them).
 * Citations are relative to /root/reference/src/.
 */
#include "aflak_oracle.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <float.h>
#ifdef _OPENMP
#include <omp.h>
#endif

static __thread char g_err[512];

const char *orc_last_error(void) { return g_err; }

int orc_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* Rust `f as usize` / `f as u8`: truncate toward zero, saturate, NaN -> 0. */
static inline uint64_t rust_f32_as_usize(float f) {
    if (!(f == f)) return 0;
    if (f <= 0.0f) return 0;
    if (f >= 18446744073709551616.0f) return UINT64_MAX;
    return (uint64_t)f;
}
static inline uint8_t rust_f32_as_u8(float f) {
    if (!(f == f)) return 0;
    if (f <= 0.0f) return 0;
    if (f >= 255.0f) return 255;
    return (uint8_t)f;
}

/* ------------------------------------------------------------------------- */
/* Preconditions: aflak_primitives/src/precond.rs                              */
/* ------------------------------------------------------------------------- */

/* try_into_unsigned! (precond.rs:3-15) */
static int check_unsigned(const char *name, int64_t v) {
    if (v >= 0) return ORC_OK;
    snprintf(g_err, sizeof g_err, "'%s' must be positive, but got %lld", name, (long long)v);
    return ORC_ERR_UNEXPECTED_INPUT;
}

/* is_sliceable!(image, start, end) (precond.rs:87-98): start < end, end < frame_cnt + 1 */
static int check_sliceable_range(const char *img, int64_t nz, int64_t start, int64_t end) {
    if (!(start < end)) {
        snprintf(g_err, sizeof g_err, "Expected start < end, but got %lld < %lld", (long long)start, (long long)end);
        return ORC_ERR_UNEXPECTED_INPUT;
    }
    if (!(end < nz + 1)) {
        snprintf(g_err, sizeof g_err,
                 "'end' greater or equal to the input image '%s''s frame count (expected %lld < %lld)", img,
                 (long long)end, (long long)(nz + 1));
        return ORC_ERR_UNEXPECTED_INPUT;
    }
    return ORC_OK;
}

/* is_sliceable!(image, frame_idx) (precond.rs:78-86) */
static int check_sliceable_frame(const char *img, const char *name, int64_t nz, int64_t frame) {
    if (!(frame < nz)) {
        snprintf(g_err, sizeof g_err,
                 "'%s' greater or equal to the input image '%s''s frame count (expected %lld < %lld)", name, img,
                 (long long)frame, (long long)nz);
        return ORC_ERR_UNEXPECTED_INPUT;
    }
    return ORC_OK;
}

static int check_reduce_args(int64_t nz, int64_t start, int64_t end) {
    int rc;
    if ((rc = check_unsigned("start", start))) return rc;
    if ((rc = check_unsigned("end", end))) return rc;
    return check_sliceable_range("image", nz, start, end);
}

/* ------------------------------------------------------------------------- */
/* Axis-0 reductions: lib.rs:1303-1330 over ndarray 0.12.1 sum_axis/mean_axis  */
/* ------------------------------------------------------------------------- */

/*
 * ndarray 0.12.1 `sum_axis(Axis(0))`, general branch:
 *   res = frame[0].to_owned(); for i in 1..n { res = res + &frame[i] }
 * i.e. per pixel a left-to-right f32 sum in k order.  Written here frame by
 * frame exactly like the reference (three memory streams per frame).
 */
int orc_integral(const float *cube, int64_t nz, int64_t n2, int64_t start, int64_t end, float *out) {
    int rc = check_reduce_args(nz, start, end);
    if (rc) return rc;
    memcpy(out, cube + start * n2, (size_t)n2 * sizeof(float));
    for (int64_t k = start + 1; k < end; ++k) {
        const float *f = cube + k * n2;
        for (int64_t i = 0; i < n2; ++i) out[i] = out[i] + f[i];
    }
    return ORC_OK;
}

/* mean_axis: sum_axis / cnt, cnt = 1.0 + 1.0 + ... (n-1 additions), true division */
static float mean_count(int64_t n) {
    float cnt = 1.0f;
    for (int64_t i = 1; i < n; ++i) cnt = cnt + 1.0f;
    return cnt;
}

int orc_average(const float *cube, int64_t nz, int64_t n2, int64_t start, int64_t end, float *out) {
    int rc = orc_integral(cube, nz, n2, start, end, out);
    if (rc) return rc;
    float cnt = mean_count(end - start);
    for (int64_t i = 0; i < n2; ++i) out[i] = out[i] / cnt;
    return ORC_OK;
}

/*
 * ndarray 0.12.1 numeric_util::unrolled_fold, reached by sum_axis only for a
 * 2-D input whose summed axis has stride 1 (shape [nz,1]).
 */
int orc_integral_lane(const float *col, int64_t nz, int64_t start, int64_t end, float *out) {
    int rc = check_reduce_args(nz, start, end);
    if (rc) return rc;
    const float *xs = col + start;
    int64_t len = end - start;
    float acc = 0.0f, p[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    while (len >= 8) {
        for (int j = 0; j < 8; ++j) p[j] = p[j] + xs[j];
        xs += 8;
        len -= 8;
    }
    acc = acc + (p[0] + p[4]);
    acc = acc + (p[1] + p[5]);
    acc = acc + (p[2] + p[6]);
    acc = acc + (p[3] + p[7]);
    for (int64_t i = 0; i < len && i < 7; ++i) acc = acc + xs[i];
    *out = acc;
    return ORC_OK;
}

void orc_sum_f64(const float *cube, int64_t n2, int64_t start, int64_t end, double *sum, double *abs_sum) {
    for (int64_t i = 0; i < n2; ++i) {
        sum[i] = 0.0;
        if (abs_sum) abs_sum[i] = 0.0;
    }
    for (int64_t k = start; k < end; ++k) {
        const float *f = cube + k * n2;
        for (int64_t i = 0; i < n2; ++i) {
            sum[i] += (double)f[i];
            if (abs_sum) abs_sum[i] += fabs((double)f[i]);
        }
    }
}

/* ------------------------------------------------------------------------- */
/* FITS data unit -> f32: unit.rs:283-331, fits.rs:42-103 over fitrs 0.5.0     */
/* ------------------------------------------------------------------------- */

/*
 * fitrs 0.5.0 `Hdu::read_data` (external, not on disk; restated from the FITS standard): the data unit is big-endian;
 * BITPIX -32 -> FloatingPoint32, -64 -> FloatingPoint64, 16 and 32 -> IntegersI32(Option<i32>) with BLANK -> None,
 * unsigned 32 -> IntegersU32.  aflak then maps to f32 (fits.rs:42-103: `f as f32`, `int as f32`, None -> NaN) and
 * applies `image * bscale + bzero` (unit.rs:328): two separately rounded f32 steps, always executed.
 * fmt: 0 f32, 1 f64, 2 i16, 3 i32, 4 u32.
 */
int orc_fits_decode(const uint8_t *raw, int fmt, int64_t n, int has_blank, int64_t blank, float bscale, float bzero, float *out) {
    for (int64_t i = 0; i < n; ++i) {
        float v;
        switch (fmt) {
        case 0: {
            const uint8_t *p = raw + 4 * i;
            uint32_t b = ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3];
            memcpy(&v, &b, 4);
            break;
        }
        case 1: {
            const uint8_t *p = raw + 8 * i;
            uint64_t b = 0;
            for (int j = 0; j < 8; ++j) b = (b << 8) | p[j];
            double d;
            memcpy(&d, &b, 8);
            v = (float)d;
            break;
        }
        case 2: {
            const uint8_t *p = raw + 2 * i;
            int32_t x = (int16_t)(uint16_t)(((uint16_t)p[0] << 8) | p[1]);
            v = (has_blank && (int64_t)x == blank) ? NAN : (float)x;
            break;
        }
        case 3: case 4: {
            const uint8_t *p = raw + 4 * i;
            uint32_t b = ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3];
            if (fmt == 3) {
                int32_t x = (int32_t)b;
                v = (has_blank && (int64_t)x == blank) ? NAN : (float)x;
            } else {
                v = (has_blank && (int64_t)b == blank) ? NAN : (float)b;
            }
            break;
        }
        default:
            snprintf(g_err, sizeof g_err, "Unsupported data array format");
            return ORC_ERR_UNEXPECTED_INPUT;
        }
        out[i] = v * bscale + bzero;
    }
    return ORC_OK;
}

/* ------------------------------------------------------------------------- */
/* Other axis-0 reductions: lib.rs:1332-1582                                   */
/* ------------------------------------------------------------------------- */

/*
 * ndarray 0.12.1 `var_axis(Axis(0), ddof)` (called with ddof = 1.0, lib.rs:1333):
 *   count = 0; mean = zeros; sum_sq = zeros;
 *   for frame in axis_iter { count += 1; for each pixel { delta = x - mean; mean = mean + delta / count;
 *                                                         sum_sq = (x - mean).mul_add(delta, sum_sq) } }
 *   panic if ddof >= count; else sum_sq / (count - ddof)
 * `std_axis` = var_axis(..).mapv(sqrt) (lib.rs:1337).  mul_add is a fused multiply-add.
 */
int orc_variance(const float *cube, int64_t nz, int64_t n2, int64_t start, int64_t end, int take_sqrt, float *out) {
    int rc = check_reduce_args(nz, start, end);
    if (rc) return rc;
    if (end - start < 2) {
        snprintf(g_err, sizeof g_err, "`ddof` needs to be strictly smaller than the length of the axis you are computing the variance/standard deviation for!");
        return ORC_ERR_PANIC;
    }
    float *mean = (float *)calloc((size_t)(n2 > 0 ? n2 : 1), sizeof(float));
    for (int64_t i = 0; i < n2; ++i) out[i] = 0.0f; /* sum_sq */
    float count = 0.0f;
    for (int64_t k = start; k < end; ++k) {
        const float *f = cube + k * n2;
        count = count + 1.0f;
        for (int64_t i = 0; i < n2; ++i) {
            float delta = f[i] - mean[i];
            mean[i] = mean[i] + delta / count;
            out[i] = fmaf(f[i] - mean[i], delta, out[i]);
        }
    }
    float dof = count - 1.0f;
    for (int64_t i = 0; i < n2; ++i) {
        out[i] = out[i] / dof;
        if (take_sqrt) out[i] = sqrtf(out[i]);
    }
    free(mean);
    return ORC_OK;
}

/* two-pass f64 variance (ddof = 1) of the f32 inputs: the tolerance oracle */
void orc_variance_f64(const float *cube, int64_t n2, int64_t start, int64_t end, double *out) {
    const double n = (double)(end - start);
    for (int64_t i = 0; i < n2; ++i) {
        double m = 0.0, v = 0.0;
        for (int64_t k = start; k < end; ++k) m += (double)cube[k * n2 + i];
        m /= n;
        for (int64_t k = start; k < end; ++k) {
            double d = (double)cube[k * n2 + i] - m;
            v += d * d;
        }
        out[i] = v / (n - 1.0);
    }
}

static int cmp_f32(const void *a, const void *b) {
    float x = *(const float *)a, y = *(const float *)b;
    return (x > y) - (x < y);
}

/* Rust's slice::sort_by is a STABLE sort (merge sort): elements that compare equal keep their order.  Under partial_cmp
 * -0.0 == +0.0, so the sign of a zero median depends on the frame order of the zeros; qsort gives no such guarantee.
 * Bottom-up merge sort taking the right run only when it is strictly smaller. */
static void stable_sort_f32(float *a, float *tmp, int64_t n) {
    for (int64_t w = 1; w < n; w *= 2) {
        for (int64_t lo = 0; lo < n; lo += 2 * w) {
            int64_t mid = lo + w < n ? lo + w : n, hi = lo + 2 * w < n ? lo + 2 * w : n;
            int64_t i = lo, j = mid, o = lo;
            while (i < mid && j < hi) tmp[o++] = (a[j] < a[i]) ? a[j++] : a[i++];
            while (i < mid) tmp[o++] = a[i++];
            while (j < hi) tmp[o++] = a[j++];
        }
        memcpy(a, tmp, (size_t)n * sizeof(float));
    }
}

/* run_median, lib.rs:1340-1372: per pixel sort (partial_cmp().unwrap() panics on NaN), middle or mean of the two middles */
int orc_median(const float *cube, int64_t nz, int64_t n2, int64_t start, int64_t end, float *out) {
    int rc = check_reduce_args(nz, start, end);
    if (rc) return rc;
    int64_t n = end - start;
    float *vals = (float *)malloc((size_t)n * 2 * sizeof(float)), *tmp = vals + n;
    for (int64_t i = 0; i < n2; ++i) {
        for (int64_t k = 0; k < n; ++k) {
            vals[k] = cube[(start + k) * n2 + i];
            if (vals[k] != vals[k] && n > 1) {
                free(vals);
                snprintf(g_err, sizeof g_err, "called `Option::unwrap()` on a `None` value");
                return ORC_ERR_PANIC;
            }
        }
        stable_sort_f32(vals, tmp, n);
        out[i] = (n % 2 == 1) ? vals[(n - 1) / 2] : (vals[n / 2] + vals[n / 2 - 1]) / 2.0f;
    }
    free(vals);
    return ORC_OK;
}

/* run_minmax, lib.rs:1374-1398: fold_axis(Axis(0), -inf, |x, y| if x > y {x} else {y}) (max) / (+inf, x < y) (min) */
int orc_minmax_fold(const float *cube, int64_t nz, int64_t n2, int64_t start, int64_t end, int is_min, float *out) {
    int rc = check_reduce_args(nz, start, end);
    if (rc) return rc;
    for (int64_t i = 0; i < n2; ++i) {
        float acc = is_min ? INFINITY : -INFINITY;
        for (int64_t k = start; k < end; ++k) {
            float y = cube[k * n2 + i];
            if (is_min) acc = (acc < y) ? acc : y;
            else acc = (acc > y) ? acc : y;
        }
        out[i] = acc;
    }
    return ORC_OK;
}

/*
 * WcsArray::pix2world (unit.rs:335-342) over fitrs 0.5.0 `WCS` (external, not on disk): restated as the
 * axis-aligned linear transform world = CRVAL + CDELT * (pixel + 1 - CRPIX) in f64, cast to f32.  wcs = {crval,
 * crpix, cdelt} or NULL (no WCS keys: the pixel maps to itself).  Parity unpinned.
 */
float orc_pix2world(const double *wcs, float pixel) {
    if (!wcs) return pixel;
    return (float)(wcs[0] + wcs[2] * ((double)pixel + (1.0 - wcs[1])));
}

/*
 * run_create_argmap, lib.rs:1400-1460: strict comparison from -inf (max) / +inf (min), `out` updated whenever a
 * frame wins: out = pix2world(trypix, ((k + start) as i64 + range) as f32) when is_actual_value (wcs != NULL or
 * identity), else the index itself.  Nothing won (all NaN) -> 0.0.
 */
int orc_argmap(const float *cube, int64_t nz, int64_t n2, int64_t start, int64_t end, int64_t range, int is_min,
               const double *wcs, float *out) {
    int rc = check_reduce_args(nz, start, end);
    if (rc) return rc;
    for (int64_t i = 0; i < n2; ++i) {
        float value = is_min ? INFINITY : -INFINITY;
        float o = 0.0f;
        for (int64_t k = 0; k < end - start; ++k) {
            float x = cube[(start + k) * n2 + i];
            if ((!is_min && x > value) || (is_min && x < value)) {
                value = x;
                o = orc_pix2world(wcs, (float)((k + start) + range));
            }
        }
        out[i] = o;
    }
    return ORC_OK;
}

/* run_centroid, lib.rs:1466-1527: flux_sum and sum(flux * wavelength) as sequential f32 sums from 0.0, then waveimg / flux_sum */
int orc_centroid(const float *cube, int64_t nz, int64_t n2, int64_t start, int64_t end, const double *wcs, float *out) {
    int rc = check_reduce_args(nz, start, end);
    if (rc) return rc;
    for (int64_t i = 0; i < n2; ++i) {
        float fs = 0.0f, ws = 0.0f;
        for (int64_t k = 0; k < end - start; ++k) {
            float flux = cube[(start + k) * n2 + i];
            float wavelength = orc_pix2world(wcs, (float)(k + start));
            fs += flux;
            ws += flux * wavelength;
        }
        out[i] = ws / fs;
    }
    return ORC_OK;
}

/* f64 version of the same quotient (products of the f32 inputs and f32 wavelengths): the tolerance oracle */
void orc_centroid_f64(const float *cube, int64_t n2, int64_t start, int64_t end, const double *wcs, double *out) {
    for (int64_t i = 0; i < n2; ++i) {
        double fs = 0.0, ws = 0.0;
        for (int64_t k = start; k < end; ++k) {
            double flux = (double)cube[k * n2 + i];
            fs += flux;
            ws += flux * (double)orc_pix2world(wcs, (float)k);
        }
        out[i] = ws / fs;
    }
}

/* run_centroid_with_mask, lib.rs:1529-1582: per-pixel [start_mask as usize, end_mask as usize); slice_axis panics when
 * start > end or end > nz; wavelength from pix2world(2, k) */
int orc_centroid_mask(const float *cube, int64_t nz, int64_t n2, const float *start_mask, const float *end_mask,
                      const double *wcs, float *out) {
    for (int64_t i = 0; i < n2; ++i) {
        uint64_t lo = rust_f32_as_usize(start_mask[i]), hi = rust_f32_as_usize(end_mask[i]);
        if (lo > hi || hi > (uint64_t)nz) {
            snprintf(g_err, sizeof g_err, "slice index out of range");
            return ORC_ERR_PANIC;
        }
        float fs = 0.0f, ws = 0.0f;
        for (uint64_t k = lo; k < hi; ++k) {
            float flux = cube[(int64_t)k * n2 + i];
            fs += flux;
            ws += flux * orc_pix2world(wcs, (float)k);
        }
        out[i] = ws / fs;
    }
    return ORC_OK;
}

/* lib.rs:393-395: middle = (start as f32 + end as f32) / 2.0; width = end as f32 - start as f32 */
void orc_middle_width(int64_t start, int64_t end, float *middle, float *width) {
    *middle = ((float)start + (float)end) / 2.0f;
    *width = (float)end - (float)start;
}

/*
 * All-cores baseline: same per-pixel left-to-right order, pixels partitioned
 * into contiguous tiles across threads (a generous reading of "all the host
 * threads it can use"; the reference's sum_axis itself is single-threaded).
 */
int orc_integral_mt(const float *cube, int64_t nz, int64_t n2, int64_t start, int64_t end, float *out, int threads) {
    int rc = check_reduce_args(nz, start, end);
    if (rc) return rc;
    if (threads < 1) threads = 1;
    const int64_t tile = 4096; /* 16 KiB of accumulators stays in L1 */
    int64_t ntiles = (n2 + tile - 1) / tile;
#pragma omp parallel for schedule(dynamic, 1) num_threads(threads)
    for (int64_t t = 0; t < ntiles; ++t) {
        int64_t i0 = t * tile, i1 = i0 + tile < n2 ? i0 + tile : n2;
        float *o = out + i0;
        int64_t w = i1 - i0;
        memcpy(o, cube + start * n2 + i0, (size_t)w * sizeof(float));
        for (int64_t k = start + 1; k < end; ++k) {
            const float *f = cube + k * n2 + i0;
            for (int64_t i = 0; i < w; ++i) o[i] = o[i] + f[i];
        }
    }
    return ORC_OK;
}

int orc_average_mt(const float *cube, int64_t nz, int64_t n2, int64_t start, int64_t end, float *out, int threads) {
    int rc = orc_integral_mt(cube, nz, n2, start, end, out, threads);
    if (rc) return rc;
    float cnt = mean_count(end - start);
    if (threads < 1) threads = 1;
#pragma omp parallel for schedule(static) num_threads(threads)
    for (int64_t i = 0; i < n2; ++i) out[i] = out[i] / cnt;
    return ORC_OK;
}

/* ------------------------------------------------------------------------- */
/* Slicing / gather                                                           */
/* ------------------------------------------------------------------------- */

/* run_slice_one_frame, lib.rs:912-925: index_axis(Axis(0), frame).to_owned() */
int orc_slice_one_frame(const float *cube, int64_t nz, int64_t n2, int64_t frame, float *out) {
    int rc;
    if ((rc = check_unsigned("frame_idx", frame))) return rc;
    if ((rc = check_sliceable_frame("input_img", "frame_idx", nz, frame))) return rc;
    memcpy(out, cube + frame * n2, (size_t)n2 * sizeof(float));
    return ORC_OK;
}

/* run_down_sampling, lib.rs:1894-1962: decimation, keep (x % n == 0 && y % n == 0) */
int orc_down_sampling(const float *img, int ndim, const int64_t *shape, int64_t n, float *out, int64_t *out_shape) {
    if (!(n > 0)) {
        snprintf(g_err, sizeof g_err, "n should be greater than 0, %lld > 0", (long long)n);
        return ORC_ERR_UNEXPECTED_INPUT;
    }
    if (ndim != 2 && ndim != 3) {
        snprintf(g_err, sizeof g_err, "This is neither 2 nor 3 dimensional image!");
        return ORC_ERR_UNEXPECTED_INPUT;
    }
    int64_t c = ndim == 3 ? shape[0] : 1;
    int64_t ny = shape[ndim - 2], nx = shape[ndim - 1];
    int64_t oy = ny / n + (ny % n != 0), ox = nx / n + (nx % n != 0);
    int64_t w = 0;
    for (int64_t z = 0; z < c; ++z) {
        const float *f = img + z * ny * nx;
        int64_t counter = 0;
        for (int64_t i = 0; i < ny * nx; ++i) {
            int64_t cx = counter % nx, cy = counter / nx;
            if (cx % n == 0 && cy % n == 0) out[w++] = f[i];
            counter += 1;
        }
    }
    if (ndim == 3) {
        out_shape[0] = c;
        out_shape[1] = oy;
        out_shape[2] = ox;
    } else {
        out_shape[0] = oy;
        out_shape[1] = ox;
    }
    return ORC_OK;
}

/* ------------------------------------------------------------------------- */
/* Elementwise operators                                                      */
/* ------------------------------------------------------------------------- */

/* run_linear_composition, lib.rs:1288-1297: i1 * coef1 + i2 * coef2, three rounded steps */
void orc_linear_composition(const float *u, const float *v, float a, float b, int64_t n, float *out) {
    for (int64_t i = 0; i < n; ++i) {
        float ua = u[i] * a;
        float vb = v[i] * b;
        out[i] = ua + vb;
    }
}

/* &WcsArray / &WcsArray, unit.rs:863-880 */
void orc_ratio(const float *u, const float *v, int64_t n, float *out) {
    for (int64_t i = 0; i < n; ++i) out[i] = u[i] / v[i];
}

/* run_image_multiplier, lib.rs:1619-1638: powf(u,a) * powf(v,b) */
void orc_image_multiplier(const float *u, const float *v, float a, float b, int64_t n, float *out) {
    for (int64_t i = 0; i < n; ++i) {
        float pu = powf(u[i], a);
        float pv = powf(v[i], b);
        out[i] = pu * pv;
    }
}

/* run_create_equivalent_width, lib.rs:1584-1596:
 *   t = (i_off - i_on) * fl / i_off * (is_emission ? -1 : 1); v > max -> 0 */
void orc_equivalent_width(const float *off, const float *on, float fl, float max, int is_emission, int64_t n, float *out) {
    float s = is_emission ? -1.0f : 1.0f;
    for (int64_t i = 0; i < n; ++i) {
        float t = off[i] - on[i];
        t = t * fl;
        t = t / off[i];
        t = t * s;
        out[i] = (t > max) ? 0.0f : t;
    }
}

/* image_min_max, lib.rs:180-197: Rust f32::min/max ignore a NaN operand */
void orc_image_min_max(const float *img, int64_t n, float *vmin, float *vmax) {
    float mn = FLT_MAX, mx = -FLT_MAX;
    for (int64_t i = 0; i < n; ++i) {
        mn = fminf(mn, img[i]);
        mx = fmaxf(mx, img[i]);
    }
    *vmin = mn;
    *vmax = mx;
}

/* run_convert_to_logscale, lib.rs:1598-1611 */
static inline float logscale_one(float v, float a, float vmin, float vmax, float ln_a) {
    float d = (v - vmin) / (vmax - vmin);
    float t = a * d;
    t = t + 1.0f;
    return logf(t) / ln_a;
}

void orc_convert_to_logscale(const float *img, int64_t n, float a, float vmin, float vmax, float *out) {
    float ln_a = logf(a);
    for (int64_t i = 0; i < n; ++i) out[i] = logscale_one(img[i], a, vmin, vmax, ln_a);
}

void orc_convert_to_logscale_mt(const float *img, int64_t n, float a, float vmin, float vmax, float *out, int threads) {
    float ln_a = logf(a);
    if (threads < 1) threads = 1;
#pragma omp parallel for schedule(static) num_threads(threads)
    for (int64_t i = 0; i < n; ++i) out[i] = logscale_one(img[i], a, vmin, vmax, ln_a);
}

/* run_log10, lib.rs:1613-1617 */
void orc_log10(const float *img, int64_t n, float *out) {
    for (int64_t i = 0; i < n; ++i) out[i] = log10f(img[i]);
}

/* run_apply_tone_curve, lib.rs:1858-1871:
 *   key = (v * table_size as f32 / 65535.0) as usize; value = 65535.0 * table[key] / table_size as f32 */
int orc_apply_tone_curve(const float *img, int64_t n, const float *table, int64_t table_len, float *out) {
    if (table_len < 1) {
        snprintf(g_err, sizeof g_err, "tone curve table is empty (reference panics: attempt to subtract with overflow)");
        return ORC_ERR_PANIC;
    }
    float ts = (float)(table_len - 1);
    for (int64_t i = 0; i < n; ++i) {
        float t = img[i] * ts;
        t = t / 65535.0f;
        uint64_t key = rust_f32_as_usize(t);
        if (key >= (uint64_t)table_len) {
            snprintf(g_err, sizeof g_err, "index out of bounds: the len is %lld but the index is %llu",
                     (long long)table_len, (unsigned long long)key);
            return ORC_ERR_PANIC;
        }
        float w = 65535.0f * table[key];
        out[i] = w / ts;
    }
    return ORC_OK;
}

/* ratio_from_bands, lib.rs:226-246 */
int orc_ratio_from_bands(float z, float z1, float z2, float *off_ratio_2, float *off_ratio_1) {
    if (!(z1 < z && z < z2)) {
        snprintf(g_err, sizeof g_err, "wrong magnitude correlation (%g < %g < %g)", z1, z, z2);
        return ORC_ERR_UNEXPECTED_INPUT;
    }
    float r1 = (z - z1) / (z2 - z1);
    *off_ratio_1 = r1;
    *off_ratio_2 = 1.0f - r1;
    return ORC_OK;
}

/* ------------------------------------------------------------------------- */
/* Tone-curve table: imgui_tone_curve/src/lib.rs                               */
/* ------------------------------------------------------------------------- */

/* ToneCurveState::default, lib.rs:21-38: arr[i] = i for i in 0..256 */
int64_t orc_tone_curve_default(float *table, int64_t cap) {
    if (cap < 256) return -1;
    for (int i = 0; i < 256; ++i) table[i] = (float)i;
    return 256;
}

static int cmp_cp(const void *a, const void *b) {
    float xa = ((const float *)a)[0], xb = ((const float *)b)[0];
    return (xa > xb) - (xa < xb);
}

#define PUSH(val)                    \
    do {                             \
        if (len >= cap) return -1;   \
        table[len++] = (val);        \
    } while (0)

/* create_curve, lib.rs:68-158 */
int64_t orc_create_curve(const float *cp_xy, int64_t n_cp, float *table, int64_t cap) {
    int64_t len = 0;
    if (n_cp < 2) return 0;
    if (n_cp == 2) {
        float slope = (cp_xy[3] - cp_xy[1]) / (cp_xy[2] - cp_xy[0]);
        float ys = cp_xy[1];
        for (int i = 0; i < 256; ++i) {
            float t = (float)i / 256.0f;
            float y = t * slope;
            y = y + ys;
            PUSH(y * 256.0f);
        }
        return len;
    }
    float *cp = (float *)malloc(sizeof(float) * 2 * (size_t)n_cp);
    memcpy(cp, cp_xy, sizeof(float) * 2 * (size_t)n_cp);
    /* Rust sort_by is stable; control points with equal x are not expected */
    qsort(cp, (size_t)n_cp, 2 * sizeof(float), cmp_cp);
#define CPX(i) cp[2 * (i)]
#define CPY(i) cp[2 * (i) + 1]
    if (n_cp == 3) {
        float x1 = CPX(0), x2 = CPX(1), x3 = CPX(2), y1 = CPY(0), y2 = CPY(1), y3 = CPY(2);
        float num = (y1 - y2) * (x1 - x3);
        float num2 = (y1 - y3) * (x1 - x2);
        float den = (x1 - x2) * (x1 - x3);
        den = den * (x2 - x3);
        float a = (num - num2) / den;
        float b0 = (y1 - y2) / (x1 - x2);
        float b1 = a * (x1 + x2);
        float b = b0 - b1;
        float c0 = a * x1;
        c0 = c0 * x1;
        float c1 = b * x1;
        float c = y1 - c0;
        c = c - c1;
        for (int i = 0; i < 256; ++i) {
            float t = (float)i / 256.0f;
            float q = a * t;
            q = q * t;
            float l = b * t;
            float y = q + l;
            y = y + c;
            if (len >= cap) { free(cp); return -1; }
            table[len++] = y * 256.0f;
        }
        free(cp);
        return len;
    }
    for (int64_t counter = 0; counter < n_cp - 1; ++counter) {
        if (counter == 0) {
            float p0y = CPY(0), p1x = CPX(1), p1y = CPY(1), p2y = CPY(2);
            int64_t size = (int64_t)rust_f32_as_usize(ceilf(p1x * 256.0f));
            float by = p0y - 2.0f * p1y;
            by = by + p2y;
            float cy = -3.0f * p0y + 4.0f * p1y;
            cy = cy - p2y;
            float dy = 2.0f * p0y;
            for (int64_t i = 0; i < size; ++i) {
                float t = (float)i / (float)(size - 1);
                float q = by * t;
                q = q * t;
                float l = cy * t;
                float y = q + l;
                y = y + dy;
                if (len >= cap) { free(cp); return -1; }
                table[len++] = y * 128.0f;
            }
        } else if (counter == n_cp - 2) {
            float p0y = CPY(counter - 1), p1x = CPX(counter), p1y = CPY(counter), p2x = CPX(counter + 1),
                  p2y = CPY(counter + 1);
            int64_t size = (int64_t)rust_f32_as_usize(floorf(p2x * 256.0f) - ceilf(p1x * 256.0f));
            float by = p0y - 2.0f * p1y;
            by = by + p2y;
            float cy = -p0y + p2y;
            float dy = 2.0f * p1y;
            for (int64_t i = 0; i < size; ++i) {
                float t = (float)i / (float)(size - 1);
                float q = by * t;
                q = q * t;
                float l = cy * t;
                float y = q + l;
                y = y + dy;
                if (len >= cap) { free(cp); return -1; }
                table[len++] = y * 128.0f;
            }
        } else {
            float p0y = CPY(counter - 1), p1x = CPX(counter), p1y = CPY(counter), p2x = CPX(counter + 1),
                  p2y = CPY(counter + 1), p3y = CPY(counter + 2);
            int64_t size = (int64_t)rust_f32_as_usize(floorf(p2x * 256.0f) - ceilf(p1x * 256.0f)) + 1;
            float ay = -p0y + 3.0f * p1y;
            ay = ay - 3.0f * p2y;
            ay = ay + p3y;
            float by = 2.0f * p0y - 5.0f * p1y;
            by = by + 4.0f * p2y;
            by = by - p3y;
            float cy = -p0y + p2y;
            float dy = 2.0f * p1y;
            for (int64_t i = 0; i < size; ++i) {
                float t = (float)i / (float)(size - 1);
                float c3 = ay * t;
                c3 = c3 * t;
                c3 = c3 * t;
                float q = by * t;
                q = q * t;
                float l = cy * t;
                float y = c3 + q;
                y = y + l;
                y = y + dy;
                if (len >= cap) { free(cp); return -1; }
                table[len++] = y * 128.0f;
            }
        }
    }
    free(cp);
    return len;
}

/* ------------------------------------------------------------------------- */
/* Colour LUT: aflak_plot/src/imshow/lut.rs                                    */
/* ------------------------------------------------------------------------- */

/* StopIter, lut.rs:251-286: (0, c_first), (lims0 + (lims2-lims0)*v_i, c_i)..., (1, c_last) */
static int64_t stop_iter(const float *stop_v, const uint8_t *stop_rgb, int64_t n, const float lims[3], int64_t i,
                         float *v, uint8_t c[3]) {
    if (n == 0) return 0;
    if (i == 0) {
        *v = 0.0f;
        memcpy(c, stop_rgb, 3);
        return 1;
    }
    if (i - 1 == n) {
        *v = 1.0f;
        memcpy(c, stop_rgb + 3 * (n - 1), 3);
        return 1;
    }
    if (i - 1 < n) {
        float span = lims[2] - lims[0];
        float t = span * stop_v[i - 1];
        *v = lims[0] + t;
        memcpy(c, stop_rgb + 3 * (i - 1), 3);
        return 1;
    }
    return 0;
}

int64_t orc_lut_bounds(const float *stop_v, const uint8_t *stop_rgb, int64_t n_stops, const float lims[3], float *v1,
                       uint8_t *c1, float *v2, uint8_t *c2, int64_t cap) {
    int64_t cnt = 0;
    for (int64_t i = 0;; ++i) {
        float a, b;
        uint8_t ca[3], cb[3];
        if (!stop_iter(stop_v, stop_rgb, n_stops, lims, i, &a, ca)) break;
        if (!stop_iter(stop_v, stop_rgb, n_stops, lims, i + 1, &b, cb)) break;
        if (cnt < cap) {
            v1[cnt] = a;
            v2[cnt] = b;
            memcpy(c1 + 3 * cnt, ca, 3);
            memcpy(c2 + 3 * cnt, cb, 3);
        }
        cnt++;
    }
    return cnt;
}

/* color_at_init, lut.rs:162-188 */
static void color_at_init(const float *stop_v, const uint8_t *stop_rgb, int64_t n, const float lims[3], float point,
                          uint8_t out[3]) {
    for (int64_t i = 0;; ++i) {
        float v1, v2;
        uint8_t c1[3], c2[3];
        if (!stop_iter(stop_v, stop_rgb, n, lims, i, &v1, c1)) break;
        if (!stop_iter(stop_v, stop_rgb, n, lims, i + 1, &v2, c2)) break;
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
                    out[ch] = rust_f32_as_u8(a + d);
                }
            }
            return;
        }
    }
    out[0] = out[1] = out[2] = 0;
}

/* lut_init, lut.rs:190-194 */
void orc_lut_build(const float *stop_v, const uint8_t *stop_rgb, int64_t n_stops, const float lims[3], uint8_t *lut) {
    for (int i = 0; i < ORC_LUT_SIZE; ++i) {
        float p = (float)i / (float)(ORC_LUT_SIZE - 1);
        color_at_init(stop_v, stop_rgb, n_stops, lims, p, lut + 3 * i);
    }
}

/* mtf, lut.rs:152-160 */
float orc_mtf(float m, float x) {
    if (x >= 1.0f) return 1.0f;
    if (x <= 0.0f) return 0.0f;
    float num = (m - 1.0f) * x;
    float den = 2.0f * m;
    den = den - 1.0f;
    den = den * x;
    den = den - m;
    return num / den;
}

/* color_at, lut.rs:137-149 (index part) */
uint32_t orc_lut_index(float point, const float lims[3]) {
    float c = (point - lims[0]) / (lims[2] - lims[0]);
    float i = orc_mtf(lims[1], c) * (float)(ORC_LUT_SIZE - 1);
    if (i < 0.0f) i = 0.0f;
    uint64_t idx = rust_f32_as_usize(i);
    if (idx >= ORC_LUT_SIZE) idx = ORC_LUT_SIZE - 1;
    return (uint32_t)idx;
}

void orc_color_at(const uint8_t *lut, const float lims[3], float point, uint8_t rgb[3]) {
    uint32_t i = orc_lut_index(point, lims);
    memcpy(rgb, lut + 3 * i, 3);
}

/* util::clamp, aflak_plot/src/util.rs:3-14 (NaN passes through both comparisons) */
static inline float clampf(float v, float mn, float mx) {
    if (v < mn) return mn;
    if (v > mx) return mx;
    return v;
}

/* color_at_bounds, lut.rs:132-135 */
void orc_color_at_bounds(const uint8_t *lut, const float lims[3], float point, float vmin, float vmax, uint8_t rgb[3]) {
    float p = clampf(point, vmin, vmax);
    orc_color_at(lut, lims, (p - vmin) / (vmax - vmin), rgb);
}

/*
 * get_vmin, lims.rs:7-20 with float_compare_nan_max (:116-124): Iterator::min_by
 * keeps the FIRST of equal elements; NaN compares greater than any number.
 */
int orc_get_vmin(const float *img, int64_t n, float *vmin) {
    if (n <= 0) return 1;
    float best = img[0];
    for (int64_t i = 1; i < n; ++i) {
        float f = img[i];
        /* replace when compare(best, f) == Greater */
        int greater;
        if (best == best && f == f) greater = best > f;
        else greater = (best != best) && (f == f);
        if (greater) best = f;
    }
    *vmin = best;
    return 0;
}

/* get_vmax, lims.rs:22-35 with float_compare_nan_min (:108-115): max_by keeps the LAST of equal elements */
int orc_get_vmax(const float *img, int64_t n, float *vmax) {
    if (n <= 0) return 1;
    float best = img[0];
    for (int64_t i = 1; i < n; ++i) {
        float f = img[i];
        /* replace unless compare(best, f) == Greater */
        int greater;
        if (best == best && f == f) greater = best > f;
        else greater = (best == best) && (f != f);
        if (!greater) best = f;
    }
    *vmax = best;
    return 0;
}

static inline void put_px(uint8_t *dst, int channels, const uint8_t rgb[3]) {
    dst[0] = rgb[0];
    dst[1] = rgb[1];
    dst[2] = rgb[2];
    if (channels == 4) dst[3] = 255;
}

/* make_raw_image, image.rs:20-49, on image.slice(s![..;-1, ..]) (image.rs:123-129) */
int orc_make_raw_image(const float *img, int64_t ny, int64_t nx, float vmin, float vmax, const uint8_t *lut,
                       const float lims[3], int channels, uint8_t *out) {
    if (vmin != vmin || vmax != vmax) {
        snprintf(g_err, sizeof g_err, "vmin, vmax not set");
        return ORC_ERR_UNEXPECTED_INPUT;
    }
    for (int64_t y = 0; y < ny; ++y) {
        const float *row = img + (ny - 1 - y) * nx;
        uint8_t *dst = out + y * nx * channels;
        for (int64_t x = 0; x < nx; ++x) {
            uint8_t rgb[3];
            orc_color_at_bounds(lut, lims, row[x], vmin, vmax, rgb);
            put_px(dst + x * channels, channels, rgb);
        }
    }
    return ORC_OK;
}

int orc_make_raw_image_mt(const float *img, int64_t ny, int64_t nx, float vmin, float vmax, const uint8_t *lut,
                          const float lims[3], int channels, uint8_t *out, int threads) {
    if (vmin != vmin || vmax != vmax) {
        snprintf(g_err, sizeof g_err, "vmin, vmax not set");
        return ORC_ERR_UNEXPECTED_INPUT;
    }
    if (threads < 1) threads = 1;
#pragma omp parallel for schedule(static) num_threads(threads)
    for (int64_t y = 0; y < ny; ++y) {
        const float *row = img + (ny - 1 - y) * nx;
        uint8_t *dst = out + y * nx * channels;
        for (int64_t x = 0; x < nx; ++x) {
            uint8_t rgb[3];
            orc_color_at_bounds(lut, lims, row[x], vmin, vmax, rgb);
            put_px(dst + x * channels, channels, rgb);
        }
    }
    return ORC_OK;
}

/* make_raw_image_rgb, image.rs:51-84 (its NaN guard `== NAN` never fires, :62-63).
 * coerce_to_array_view3 flips axis 1 (rows) the same way. */
int orc_make_raw_image_rgb(const float *img, int64_t ny, int64_t nx, const float vmin[3], const float vmax[3],
                           const uint8_t *lut3, const float *lims3, int channels, uint8_t *out) {
    int64_t n2 = ny * nx;
    for (int64_t y = 0; y < ny; ++y) {
        for (int64_t x = 0; x < nx; ++x) {
            uint8_t px[3];
            for (int c = 0; c < 3; ++c) {
                uint8_t rgb[3];
                float v = img[c * n2 + (ny - 1 - y) * nx + x];
                orc_color_at_bounds(lut3 + (size_t)c * ORC_LUT_SIZE * 3, lims3 + 3 * c, v, vmin[c], vmax[c], rgb);
                px[c] = rgb[c];
            }
            put_px(out + (y * nx + x) * channels, channels, px);
        }
    }
    return ORC_OK;
}

/* ------------------------------------------------------------------------- */
/* Spatial reductions / gathers and the remaining elementwise operators        */
/* ------------------------------------------------------------------------- */

/*
 * ROI::filter (roi.rs:15-37) followed by the sum of run_extract_wave (lib.rs:1147-1169): res = 0.0; res += val.
 * ROI::All iterates `for j in 0..size.1 { for i in 0..size.0 { data.get([j, i]).unwrap() } }` with size = (ny, nx):
 * the bounds are swapped, so a non-square frame indexes out of range and the unwrap panics.
 * PixelList: (i, j) -> data.get([j, i]), pixels outside the frame are skipped.  xy = npix pairs (i, j).
 * out64 (optional) receives the f64 sums of the same values: the tolerance oracle.
 */
int orc_extract_wave(const float *cube, int64_t nz, int64_t ny, int64_t nx, int roi_all, const uint64_t *xy, int64_t npix,
                     float *out, double *out64) {
    if (roi_all && ny != nx) {
        snprintf(g_err, sizeof g_err, "called `Option::unwrap()` on a `None` value");
        return ORC_ERR_PANIC;
    }
    for (int64_t k = 0; k < nz; ++k) {
        const float *data = cube + k * ny * nx;
        float res = 0.0f;
        double res64 = 0.0;
        if (roi_all) {
            for (int64_t j = 0; j < nx; ++j)
                for (int64_t i = 0; i < ny; ++i) {
                    res += data[j * nx + i];
                    res64 += (double)data[j * nx + i];
                }
        } else {
            for (int64_t p = 0; p < npix; ++p) {
                uint64_t i = xy[2 * p], j = xy[2 * p + 1];
                if (j < (uint64_t)ny && i < (uint64_t)nx) {
                    res += data[j * nx + i];
                    res64 += (double)data[j * nx + i];
                }
            }
        }
        out[k] = res;
        if (out64) out64[k] = res64;
    }
    return ORC_OK;
}

/* run_extrude, lib.rs:1233-1256: frames in reverse order, region in list order; the shape (nz, roi.datalen()) must
 * match the number of pushed values or from_shape_vec(..).unwrap() panics (ROI::All: datalen 0; skipped pixels) */
int orc_extrude(const float *cube, int64_t nz, int64_t ny, int64_t nx, int roi_all, const uint64_t *xy, int64_t npix, float *out) {
    int64_t pushed = 0;
    if (roi_all) {
        if (ny != nx) {
            snprintf(g_err, sizeof g_err, "called `Option::unwrap()` on a `None` value");
            return ORC_ERR_PANIC;
        }
        pushed = ny * nx;
    } else {
        for (int64_t p = 0; p < npix; ++p) pushed += (xy[2 * p + 1] < (uint64_t)ny && xy[2 * p] < (uint64_t)nx);
    }
    int64_t datalen = roi_all ? 0 : npix;
    if (nz * pushed != nz * datalen) {
        snprintf(g_err, sizeof g_err, "called `Result::unwrap()` on an `Err` value: ShapeError");
        return ORC_ERR_PANIC;
    }
    int64_t o = 0;
    for (int64_t k = nz - 1; k >= 0; --k)
        for (int64_t p = 0; p < npix; ++p) out[o++] = cube[(k * ny + (int64_t)xy[2 * p + 1]) * nx + (int64_t)xy[2 * p]];
    return ORC_OK;
}

/* run_make_plane3d, lib.rs:1130-1145: map[(j, i)] = p0 + i*dir1 + j*dir2, f32, evaluated left to right */
void orc_make_plane3d(const float p0[3], const float d1[3], const float d2[3], int64_t count1, int64_t count2, float *map) {
    for (int64_t j = 0; j < count2; ++j)
        for (int64_t i = 0; i < count1; ++i)
            for (int c = 0; c < 3; ++c) {
                float fi = (float)i, fj = (float)j;
                map[3 * (j * count1 + i) + c] = p0[c] + fi * d1[c] + fj * d2[c];
            }
}

/* Rust `{}` of an f32 for the magnitudes that occur in index errors */
static void rust_f32_display(float f, char *buf, size_t cap) {
    if (f != f) { snprintf(buf, cap, "NaN"); return; }
    if (isinf(f)) { snprintf(buf, cap, f > 0 ? "inf" : "-inf"); return; }
    if (f == floorf(f) && fabsf(f) < 1e16f) { snprintf(buf, cap, "%.0f", (double)f); return; }
    for (int prec = 1; prec < 12; ++prec) {
        snprintf(buf, cap, "%.*g", prec, (double)f);
        if (strtof(buf, NULL) == f && !strchr(buf, 'e')) return;
    }
    snprintf(buf, cap, "%.9g", (double)f);
}

/* run_slice_3d_to_2d, lib.rs:1096-1110: out = image[x as usize, y as usize, z as usize], nearest by truncation */
int orc_slice_3d_to_2d(const float *cube, int64_t d0, int64_t d1, int64_t d2, const float *map, int64_t n, float *out) {
    for (int64_t i = 0; i < n; ++i) {
        uint64_t a = rust_f32_as_usize(map[3 * i]), b = rust_f32_as_usize(map[3 * i + 1]), c = rust_f32_as_usize(map[3 * i + 2]);
        if (a >= (uint64_t)d0 || b >= (uint64_t)d1 || c >= (uint64_t)d2) {
            char x[64], y[64], z[64];
            rust_f32_display(map[3 * i], x, sizeof x);
            rust_f32_display(map[3 * i + 1], y, sizeof y);
            rust_f32_display(map[3 * i + 2], z, sizeof z);
            snprintf(g_err, sizeof g_err, "Input maps to out of bound pixel!: [%s, %s, %s]", x, y, z);
            return ORC_ERR_UNEXPECTED_INPUT;
        }
        out[i] = cube[((int64_t)a * d1 + (int64_t)b) * d2 + (int64_t)c];
    }
    return ORC_OK;
}

/* create_velocity_field_map, lib.rs:509-526: (image - w_0) * c / w_0 with c = 3e5 as f32 */
void orc_velocity_field(const float *img, int64_t n, float w0, float *out) {
    const float c = 3e5f;
    for (int64_t i = 0; i < n; ++i) out[i] = (img[i] - w0) * c / w0;
}

/* histogram_transformation, lib.rs:562-591 (one channel; _rgb, lib.rs:592-631, applies the same per channel) */
void orc_histogram_transformation(const float *img, int64_t n, const float parameters[3], float *out) {
    float min = FLT_MAX, max = -FLT_MAX;
    for (int64_t i = 0; i < n; ++i) {
        if (!(img[i] != img[i])) { /* f32::min / max ignore a NaN operand */
            min = img[i] < min ? img[i] : min;
            max = img[i] > max ? img[i] : max;
        }
    }
    for (int64_t i = 0; i < n; ++i) {
        float point = (img[i] - min) / (max - min);
        float x = (point - parameters[0]) / (parameters[2] - parameters[0]);
        float c;
        if (x > 1.0f) c = 1.0f;
        else if (x < 0.0f) c = 0.0f;
        else {
            float m = parameters[1];
            c = (m - 1.0f) * x / ((2.0f * m - 1.0f) * x - m);
        }
        out[i] = c * (max - min) + min;
    }
}

/* run_apply_arcsinh_stretch, lib.rs:1873-1892 ([3][n] image); asinh is libm::asinh (f64) in the reference, glibc's here */
void orc_arcsinh_stretch(const float *img, int64_t n, float beta, float *out) {
    for (int64_t i = 0; i < n; ++i) {
        float d0 = img[i], d1 = img[n + i], d2 = img[2 * n + i];
        float l = (d0 + d1 + d2) / 3.0f;
        double s64 = asinh((double)(l * beta)) / ((double)l * asinh((double)beta));
        float s = (float)s64;
        out[i] = s * d0;
        out[n + i] = s * d1;
        out[2 * n + i] = s * d2;
    }
}

static float rust_max(float a, float b) { return (a != a) ? b : (b != b) ? a : (a > b ? a : b); }
static float rust_min(float a, float b) { return (a != a) ? b : (b != b) ? a : (a < b ? a : b); }

/* run_color_image_to_hsv, lib.rs:1795-1826 ([3][n] image) */
void orc_color_image_to_hsv(const float *img, int64_t n, float *out) {
    for (int64_t i = 0; i < n; ++i) {
        float d0 = img[i], d1 = img[n + i], d2 = img[2 * n + i];
        float hue = 0.0f;
        float max = rust_max(rust_max(d0, d1), d2), min = rust_min(rust_min(d0, d1), d2);
        if (d0 >= d1 && d0 >= d2) hue = 60.0f * ((d1 - d2) / (max - min));
        else if (d1 >= d0 && d1 >= d2) hue = 60.0f * ((d2 - d0) / (max - min)) + 120.0f;
        else if (d2 >= d0 && d2 >= d1) hue = 60.0f * ((d0 - d1) / (max - min)) + 240.0f;
        if (hue < 0.0f) hue += 360.0f;
        hue = hue / 360.0f * 65535.0f;
        out[i] = hue;
        out[n + i] = (max - min) / max * 65535.0f;
        out[2 * n + i] = max;
    }
}

/* ------------------------------------------------------------------------- */
/* Display statistics: aflak_plot/src/imshow/hist.rs, lims.rs:37-106, zscale.rs */
/* ------------------------------------------------------------------------- */

/* hist::histogram (hist.rs:10-36) and histogram_color (hist.rs:38-66): counts is [nb][channels] */
void orc_histogram(const float *img, int64_t n_per_channel, int channels, float min, float max, int nb, uint64_t *counts) {
    for (int64_t i = 0; i < (int64_t)nb * channels; ++i) counts[i] = 0;
    for (int c = 0; c < channels; ++c)
        for (int64_t k = 0; k < n_per_channel; ++k) {
            float val = img[(int64_t)c * n_per_channel + k];
            float fi = (val - min) / (max - min) * (float)nb;
            uint64_t i = rust_f32_as_usize(fi);
            if (i >= (uint64_t)nb) i = (uint64_t)nb - 1;
            counts[i * channels + c] += 1;
        }
}

/* Bin.start / Bin.end (hist.rs:16-23) */
void orc_histogram_bins(float min, float max, int nb, float *starts, float *ends) {
    float size = (max - min) / (float)nb;
    for (int i = 0; i < nb; ++i) {
        starts[i] = min + (float)i * size;
        ends[i] = min + (float)(i + 1) * size;
    }
}

static int median_of_sorted(float *data, int64_t len, float *out) {
    if (len == 0) return 1;
    qsort(data, (size_t)len, sizeof(float), cmp_f32);
    int64_t mid = len / 2;
    *out = (len % 2 == 0) ? (data[mid] + data[mid - 1]) / 2.0f : data[mid];
    return 0;
}

/* lims::get_vmed_normalized (lims.rs:37-67); vmin/vmax from orc_get_vmin/orc_get_vmax. Returns 1 for "Empty data!, med" */
int orc_vmed_normalized(const float *img, int64_t n, float vmin, float vmax, float *out) {
    float *data = (float *)malloc((size_t)(n > 0 ? n : 1) * sizeof(float));
    int64_t len = 0;
    for (int64_t i = 0; i < n; ++i) {
        float d = (img[i] - vmin) / (vmax - vmin);
        if (!isnan(d) && !isinf(d)) data[len++] = d;
    }
    int rc = median_of_sorted(data, len, out);
    free(data);
    return rc;
}

/* lims::get_vmad_normalized (lims.rs:69-106) with the median already known. Returns 1 for "Empty data!, mad" */
int orc_vmad_normalized(const float *img, int64_t n, float vmin, float vmax, float vmed, float *out) {
    float *data = (float *)malloc((size_t)(n > 0 ? n : 1) * sizeof(float));
    int64_t len = 0;
    for (int64_t i = 0; i < n; ++i) {
        float d = fabsf((img[i] - vmin) / (vmax - vmin) - vmed);
        if (!isnan(d) && !isinf(d)) data[len++] = d;
    }
    int rc = median_of_sorted(data, len, out);
    free(data);
    return rc;
}

static int32_t rust_f32_as_i32(float f) {
    if (f != f) return 0;
    if (f >= 2147483648.0f) return INT32_MAX;
    if (f <= -2147483648.0f) return INT32_MIN;
    return (int32_t)f;
}

/* zsc_compute_sigma, zscale.rs:205-241 */
static void zsc_sigma(const float *flat, const int *badpix, int npix, int *ngoodpix, int *has_mean, int *has_sigma, float *sigma) {
    float sumz = 0.0f, sumsq = 0.0f;
    int ng = 0;
    for (int i = 0; i < npix; ++i)
        if (badpix[i] == 0) sumz = sumz + flat[i];
    for (int i = 0; i < npix; ++i)
        if (badpix[i] == 0) sumsq = sumsq + flat[i] * flat[i];
    for (int i = 0; i < npix; ++i) ng += badpix[i] == 0;
    *ngoodpix = ng;
    *has_mean = ng >= 1;
    *has_sigma = 0;
    if (ng >= 2) {
        float temp = sumsq / (float)(ng - 1) - sumz * sumz / (float)(ng * (ng - 1));
        *has_sigma = 1;
        *sigma = temp < 0.0f ? 0.0f : sqrtf(temp);
    }
}

/*
 * Image::zscale(nsamples, contrast), zscale.rs:45-100, with zsc_sample (:101-126: every stride-th pixel of every
 * stride-th row of the vertically flipped image, NaN dropped, truncated to nsamples) and zsc_fit_line (:127-203).
 * img is [nc][nl] (dim.0, dim.1).  *has = 0 is (None, None).
 */
int orc_zscale(const float *img, int64_t nc, int64_t nl, int nsamples, float contrast, int *has, float *z1, float *z2) {
    *has = 0;
    if (nc <= 0 || nl <= 0) return ORC_OK;
    uint64_t st = rust_f32_as_usize(sqrtf((float)((uint64_t)(nc - 1) * (uint64_t)(nl - 1)) / (float)nsamples));
    if (st < 1) st = 1;
    if (st > (uint64_t)INT64_MAX) st = (uint64_t)INT64_MAX;
    int64_t stride = (int64_t)st;
    int64_t cap = ((nc - 1) / stride + 1) * ((nl - 1) / stride + 1);
    float *sample = (float *)malloc((size_t)cap * sizeof(float));
    int64_t np_ = 0;
    for (int64_t r = 0; r < nc; r += stride) {
        const float *line = img + (nc - 1 - r) * nl; /* flipped rows, image.rs:123-129 */
        for (int64_t c = 0; c < nl; c += stride)
            if (!isnan(line[c])) sample[np_++] = line[c];
        if (stride > nc) break;
    }
    if (nsamples >= 0 && np_ > (int64_t)nsamples) np_ = nsamples; /* truncate(maxpix as usize) */
    if (np_ == 0) {
        free(sample);
        return ORC_OK;
    }
    qsort(sample, (size_t)np_, sizeof(float), cmp_f32);
    int npix = (int)np_;
    float zmin = sample[0], zmax = sample[npix - 1];
    int center_pixel = (npix - 1) / 2;
    float median = (npix % 2 == 1) ? sample[center_pixel] : 0.5f * (sample[center_pixel] + sample[center_pixel + 1]);
    const int MIN_NPIXELS = 5, MAX_ITERATIONS = 5;
    const float MAX_REJECT = 0.5f, KREJ = 2.5f;
    int minpix = rust_f32_as_i32((float)npix * MAX_REJECT);
    if (minpix < MIN_NPIXELS) minpix = MIN_NPIXELS;
    int ngrow = rust_f32_as_i32((float)npix * 0.01f);
    if (ngrow < 1) ngrow = 1;
    /* zsc_fit_line */
    float xscale = 2.0f / (float)(npix - 1);
    float *xnorm = (float *)malloc((size_t)npix * sizeof(float));
    float *flat = (float *)malloc((size_t)npix * sizeof(float));
    int *badpix = (int *)calloc((size_t)npix, sizeof(int));
    int *nextbad = (int *)malloc((size_t)npix * sizeof(int));
    for (int v = 0; v < npix; ++v) xnorm[v] = (float)v * xscale - 1.0f;
    int ngoodpix = npix, last_ngoodpix = npix + 1;
    float slope = 0.0f, intercept = 0.0f;
    for (int it = 0; it < MAX_ITERATIONS; ++it) {
        if (ngoodpix >= last_ngoodpix || ngoodpix < minpix) break;
        int sum = 0;
        float sumx = 0.0f, sumxx = 0.0f, sumxy = 0.0f, sumy = 0.0f;
        for (int i = 0; i < npix; ++i) if (badpix[i] == 0) { sum += 1; sumx = sumx + xnorm[i]; }
        for (int i = 0; i < npix; ++i) if (badpix[i] == 0) sumxx = sumxx + xnorm[i] * xnorm[i];
        for (int i = 0; i < npix; ++i) if (badpix[i] == 0) sumxy = sumxy + xnorm[i] * sample[i];
        for (int i = 0; i < npix; ++i) if (badpix[i] == 0) sumy = sumy + sample[i];
        float delta = (float)sum * sumxx - sumx * sumx;
        intercept = (sumxx * sumy - sumx * sumxy) / delta;
        slope = ((float)sum * sumxy - sumx * sumy) / delta;
        for (int i = 0; i < npix; ++i) {
            float fitted = xnorm[i] * slope + intercept;
            flat[i] = sample[i] - fitted;
        }
        int ng, has_mean, has_sigma;
        float sigma = 0.0f;
        zsc_sigma(flat, badpix, npix, &ng, &has_mean, &has_sigma, &sigma);
        ngoodpix = ng;
        if (has_mean && has_sigma) {
            float threshold = sigma * KREJ, lcut = -threshold, hcut = threshold;
            for (int i = 0; i < npix; ++i) nextbad[i] = (flat[i] < lcut || flat[i] > hcut) ? 1 : 0;
            /* convolve(next_badpix, vec![1, ngrow]) */
            for (int head = 0; head < npix; ++head) {
                int data = 0;
                const int filter[2] = {1, ngrow};
                for (int m = 0; m < 2; ++m) {
                    if (head < m) break;
                    data += nextbad[head - m] * filter[m];
                }
                badpix[head] = data;
            }
            ngoodpix = 0;
            for (int i = 0; i < npix; ++i) ngoodpix += badpix[i] == 0;
        }
    }
    float zslope = slope * xscale;
    (void)intercept;
    if (ngoodpix < minpix) {
        *z1 = zmin;
        *z2 = zmax;
    } else {
        if (contrast > 0.0f) zslope = zslope / contrast;
        float a = median - (float)(center_pixel - 1) * zslope, b = median + (float)(npix - center_pixel) * zslope;
        *z1 = fmaxf(zmin, a);
        *z2 = fminf(zmax, b);
    }
    *has = 1;
    free(sample); free(xnorm); free(flat); free(badpix); free(nextbad);
    return ORC_OK;
}

/* ------------------------------------------------------------------------- */
/* Synthetic generator (not from the reference; shared bit-for-bit with the    */
/* device generator in aflak_b200/csrc/gen.cuh -- only +,-,*,/ in f32)         */
/* ------------------------------------------------------------------------- */

static inline uint64_t mix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

float orc_gen_voxel(uint64_t seed, int64_t nz, int64_t ny, int64_t nx, int64_t k, int64_t y, int64_t x, int pathology) {
    uint64_t idx = ((uint64_t)k * (uint64_t)ny + (uint64_t)y) * (uint64_t)nx + (uint64_t)x;
    uint64_t h = mix64(seed + idx);
    float u1 = (float)(uint32_t)(h >> 40) * 5.9604644775390625e-08f;              /* 2^-24 */
    float u2 = (float)(uint32_t)((h >> 16) & 0xFFFFFFu) * 5.9604644775390625e-08f;
    float ph = (float)(uint32_t)((uint64_t)(x * 3 + y * 5) & 255u) * 0.00390625f; /* [0,1) */
    float fnz = (float)nz;
    float k0 = ph - 0.5f;
    k0 = k0 * 0.125f;
    k0 = k0 + 0.5f;
    k0 = k0 * fnz;
    float sigma = fnz * 0.0078125f;
    sigma = sigma + 2.0f;
    float d = (float)k - k0;
    d = d / sigma;
    float dd = d * d;
    dd = dd + 1.0f;
    float line = 2.0f / dd;
    float cont = 0.1f * u1;
    cont = cont + 1.0f;
    float noise = u2 - 0.5f;
    noise = noise * 0.05f;
    float v = cont + line;
    v = v + noise;
    if (pathology) {
        uint32_t r = (uint32_t)(h & 0xFFFFu);
        if (r < 16u) v = NAN;
        else if (r < 24u) v = INFINITY;
        else if (r < 32u) v = -INFINITY;
        else if (r < 48u) v = 0.0f;
        else if (r < 64u) v = -v;
        else if (r < 80u) v = 1e-40f;
    }
    (void)ny;
    return v;
}

void orc_gen_cube(uint64_t seed, int64_t nz, int64_t ny, int64_t nx, int64_t y0, int64_t rows, int pathology,
                  float *out, int threads) {
    if (threads < 1) threads = 1;
#pragma omp parallel for schedule(static) num_threads(threads)
    for (int64_t k = 0; k < nz; ++k) {
        for (int64_t r = 0; r < rows; ++r) {
            float *dst = out + (k * rows + r) * nx;
            for (int64_t x = 0; x < nx; ++x) dst[x] = orc_gen_voxel(seed, nz, ny, nx, k, y0 + r, x, pathology);
        }
    }
}
