"""ctypes binding of the CPU oracle (oracle/aflak_oracle.c).

TEST INFRASTRUCTURE ONLY.  Imported by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs -- never by aflak_b200/.
Parity status: LUT path pinned by the reference's golden vectors
(aflak_plot/src/imshow/lut.rs:292-343); reductions/elementwise "parity
unpinned" (no value-asserting reference test; ndarray 0.12.1 not on disk).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libaflak_oracle.so")

OK, ERR_UNEXPECTED_INPUT, ERR_PANIC = 0, 1, 2
LUT_SIZE = 65536


class OracleError(Exception):
    """IOErr::UnexpectedInput (code 1) or a reference panic (code 2)."""

    def __init__(self, code, msg):
        super().__init__(msg)
        self.code = code
        self.msg = msg


def build(force=False):
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(
        os.path.join(_HERE, "aflak_oracle.c")
    ):
        env = dict(os.environ)
        env.pop("CC", None)
        subprocess.check_call(["make", "-C", _HERE, "-s"] + (["-B"] if force else []), env=env)


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        _lib = C.CDLL(_LIB_PATH)
        _lib.orc_last_error.restype = C.c_char_p
        _lib.orc_mtf.restype = C.c_float
        _lib.orc_mtf.argtypes = [C.c_float, C.c_float]
        _lib.orc_lut_index.restype = C.c_uint32
        _lib.orc_gen_voxel.restype = C.c_float
        _lib.orc_tone_curve_default.restype = C.c_int64
        _lib.orc_create_curve.restype = C.c_int64
        _lib.orc_lut_bounds.restype = C.c_int64
    return _lib


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def _u8p(a):
    return a.ctypes.data_as(C.POINTER(C.c_uint8))


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _check(rc):
    if rc != OK:
        raise OracleError(rc, lib().orc_last_error().decode())


def max_threads():
    return int(lib().orc_max_threads())


# ---- reductions ------------------------------------------------------------

def _cube_dims(cube):
    nz = cube.shape[0] if cube.ndim > 0 else 0
    n2 = int(np.prod(cube.shape[1:])) if cube.ndim > 1 else 1
    return nz, n2


def integral(cube, start, end, threads=0):
    cube = _f32(cube)
    if cube.ndim == 0:
        raise OracleError(ERR_UNEXPECTED_INPUT, "'image' is a 0-dimensional image, cannot slice")
    nz, n2 = _cube_dims(cube)
    out = np.empty(cube.shape[1:], dtype=np.float32)
    if threads:
        _check(lib().orc_integral_mt(_fp(cube), C.c_int64(nz), C.c_int64(n2), C.c_int64(start), C.c_int64(end), _fp(out), C.c_int(threads)))
    else:
        _check(lib().orc_integral(_fp(cube), C.c_int64(nz), C.c_int64(n2), C.c_int64(start), C.c_int64(end), _fp(out)))
    return out


def average(cube, start, end, threads=0):
    cube = _f32(cube)
    if cube.ndim == 0:
        raise OracleError(ERR_UNEXPECTED_INPUT, "'image' is a 0-dimensional image, cannot slice")
    nz, n2 = _cube_dims(cube)
    out = np.empty(cube.shape[1:], dtype=np.float32)
    if threads:
        _check(lib().orc_average_mt(_fp(cube), C.c_int64(nz), C.c_int64(n2), C.c_int64(start), C.c_int64(end), _fp(out), C.c_int(threads)))
    else:
        _check(lib().orc_average(_fp(cube), C.c_int64(nz), C.c_int64(n2), C.c_int64(start), C.c_int64(end), _fp(out)))
    return out


def integral_lane(col, start, end):
    col = _f32(col).reshape(-1)
    out = np.empty(1, dtype=np.float32)
    _check(lib().orc_integral_lane(_fp(col), C.c_int64(col.size), C.c_int64(start), C.c_int64(end), _fp(out)))
    return out[0]


def sum_f64(cube, start, end):
    """f64 sum and sum of |x| over [start,end) per pixel."""
    cube = _f32(cube)
    nz, n2 = _cube_dims(cube)
    s = np.empty(cube.shape[1:], dtype=np.float64)
    a = np.empty(cube.shape[1:], dtype=np.float64)
    lib().orc_sum_f64(_fp(cube), C.c_int64(n2), C.c_int64(start), C.c_int64(end),
                      s.ctypes.data_as(C.POINTER(C.c_double)), a.ctypes.data_as(C.POINTER(C.c_double)))
    return s, a


def middle_width(start, end):
    m, w = C.c_float(), C.c_float()
    lib().orc_middle_width(C.c_int64(start), C.c_int64(end), C.byref(m), C.byref(w))
    return m.value, w.value


# ---- FITS data unit -> f32 (unit.rs:283-331, fits.rs:42-103) ----------------------

def fits_decode(raw, fmt, n, blank=None, bscale=1.0, bzero=0.0):
    """raw: bytes of the big-endian data unit; fmt: 0 f32, 1 f64, 2 i16, 3 i32, 4 u32."""
    buf = np.frombuffer(raw, dtype=np.uint8) if not isinstance(raw, np.ndarray) else raw.view(np.uint8).reshape(-1)
    out = np.empty(n, dtype=np.float32)
    _check(lib().orc_fits_decode(_u8p(buf), C.c_int(fmt), C.c_int64(n), C.c_int(0 if blank is None else 1),
                                 C.c_int64(0 if blank is None else int(blank)), C.c_float(bscale), C.c_float(bzero), _fp(out)))
    return out


# ---- other axis-0 reductions (lib.rs:1332-1582) ----------------------------------

def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _wcs(wcs):
    """wcs = (crval, crpix, cdelt) or None -> (keep-alive array, pointer)"""
    if wcs is None:
        return None, None
    w = np.asarray(wcs, dtype=np.float64)
    return w, _dp(w)


def _reduce_call(fn, cube, start, end, *extra):
    cube = _f32(cube)
    nz, n2 = _cube_dims(cube)
    out = np.empty(cube.shape[1:], dtype=np.float32)
    _check(fn(_fp(cube), C.c_int64(nz), C.c_int64(n2), C.c_int64(start), C.c_int64(end), *extra, _fp(out)))
    return out


def variance(cube, start, end):
    return _reduce_call(lib().orc_variance, cube, start, end, C.c_int(0))


def stddev(cube, start, end):
    return _reduce_call(lib().orc_variance, cube, start, end, C.c_int(1))


def variance_f64(cube, start, end):
    cube = _f32(cube)
    nz, n2 = _cube_dims(cube)
    out = np.empty(cube.shape[1:], dtype=np.float64)
    lib().orc_variance_f64(_fp(cube), C.c_int64(n2), C.c_int64(start), C.c_int64(end), _dp(out))
    return out


def median(cube, start, end):
    return _reduce_call(lib().orc_median, cube, start, end)


def minmax_fold(cube, start, end, is_min):
    return _reduce_call(lib().orc_minmax_fold, cube, start, end, C.c_int(int(is_min)))


def pix2world(wcs, pixel):
    keep, p = _wcs(wcs)
    f = lib().orc_pix2world
    f.restype = C.c_float
    return float(f(p, C.c_float(pixel)))


def argmap(cube, start, end, rng, is_min, wcs=None):
    keep, p = _wcs(wcs)
    return _reduce_call(lib().orc_argmap, cube, start, end, C.c_int64(rng), C.c_int(int(is_min)), p)


def centroid(cube, start, end, wcs=None):
    keep, p = _wcs(wcs)
    return _reduce_call(lib().orc_centroid, cube, start, end, p)


def centroid_f64(cube, start, end, wcs=None):
    cube = _f32(cube)
    nz, n2 = _cube_dims(cube)
    keep, p = _wcs(wcs)
    out = np.empty(cube.shape[1:], dtype=np.float64)
    lib().orc_centroid_f64(_fp(cube), C.c_int64(n2), C.c_int64(start), C.c_int64(end), p, _dp(out))
    return out


def centroid_mask(cube, start_mask, end_mask, wcs=None):
    cube = _f32(cube)
    nz, n2 = _cube_dims(cube)
    lo, hi = _f32(start_mask), _f32(end_mask)
    keep, p = _wcs(wcs)
    out = np.empty(cube.shape[1:], dtype=np.float32)
    _check(lib().orc_centroid_mask(_fp(cube), C.c_int64(nz), C.c_int64(n2), _fp(lo), _fp(hi), p, _fp(out)))
    return out


# ---- slicing / gather ----------------------------------------------------------

def slice_one_frame(cube, frame):
    cube = _f32(cube)
    if cube.ndim == 0:
        raise OracleError(ERR_UNEXPECTED_INPUT, "'input_img' is a 0-dimensional image, cannot slice")
    nz, n2 = _cube_dims(cube)
    out = np.empty(cube.shape[1:], dtype=np.float32)
    _check(lib().orc_slice_one_frame(_fp(cube), C.c_int64(nz), C.c_int64(n2), C.c_int64(frame), _fp(out)))
    return out


def down_sampling(img, n):
    img = _f32(img)
    shape = (C.c_int64 * max(img.ndim, 1))(*img.shape)
    oshape = (C.c_int64 * 3)()
    out = np.empty(img.size, dtype=np.float32)
    _check(lib().orc_down_sampling(_fp(img), C.c_int(img.ndim), shape, C.c_int64(n), _fp(out), oshape))
    osh = tuple(int(oshape[i]) for i in range(img.ndim))
    return out[: int(np.prod(osh))].reshape(osh).copy()


# ---- elementwise ---------------------------------------------------------------

def linear_composition(u, v, a, b):
    u, v = _f32(u), _f32(v)
    out = np.empty_like(u)
    lib().orc_linear_composition(_fp(u), _fp(v), C.c_float(a), C.c_float(b), C.c_int64(u.size), _fp(out))
    return out


def ratio(u, v):
    u, v = _f32(u), _f32(v)
    out = np.empty_like(u)
    lib().orc_ratio(_fp(u), _fp(v), C.c_int64(u.size), _fp(out))
    return out


def image_multiplier(u, v, a, b):
    u, v = _f32(u), _f32(v)
    out = np.empty_like(u)
    lib().orc_image_multiplier(_fp(u), _fp(v), C.c_float(a), C.c_float(b), C.c_int64(u.size), _fp(out))
    return out


def equivalent_width(off, on, fl, vmax, is_emission):
    off, on = _f32(off), _f32(on)
    out = np.empty_like(off)
    lib().orc_equivalent_width(_fp(off), _fp(on), C.c_float(fl), C.c_float(vmax), C.c_int(int(is_emission)), C.c_int64(off.size), _fp(out))
    return out


def image_min_max(img):
    img = _f32(img)
    mn, mx = C.c_float(), C.c_float()
    lib().orc_image_min_max(_fp(img), C.c_int64(img.size), C.byref(mn), C.byref(mx))
    return np.float32(mn.value), np.float32(mx.value)


def convert_to_logscale(img, a, vmin, vmax, threads=0):
    img = _f32(img)
    out = np.empty_like(img)
    if threads:
        lib().orc_convert_to_logscale_mt(_fp(img), C.c_int64(img.size), C.c_float(a), C.c_float(vmin), C.c_float(vmax), _fp(out), C.c_int(threads))
    else:
        lib().orc_convert_to_logscale(_fp(img), C.c_int64(img.size), C.c_float(a), C.c_float(vmin), C.c_float(vmax), _fp(out))
    return out


def log10(img):
    img = _f32(img)
    out = np.empty_like(img)
    lib().orc_log10(_fp(img), C.c_int64(img.size), _fp(out))
    return out


def apply_tone_curve(img, table):
    img, table = _f32(img), _f32(table)
    out = np.empty_like(img)
    _check(lib().orc_apply_tone_curve(_fp(img), C.c_int64(img.size), _fp(table), C.c_int64(table.size), _fp(out)))
    return out


def ratio_from_bands(z, z1, z2):
    r2, r1 = C.c_float(), C.c_float()
    _check(lib().orc_ratio_from_bands(C.c_float(z), C.c_float(z1), C.c_float(z2), C.byref(r2), C.byref(r1)))
    return np.float32(r2.value), np.float32(r1.value)


def tone_curve_default():
    t = np.empty(256, dtype=np.float32)
    n = lib().orc_tone_curve_default(_fp(t), C.c_int64(256))
    return t[:n]


def create_curve(cp):
    cp = _f32(cp).reshape(-1, 2)
    t = np.empty(4096, dtype=np.float32)
    n = lib().orc_create_curve(_fp(cp), C.c_int64(cp.shape[0]), _fp(t), C.c_int64(t.size))
    assert n >= 0
    return t[:n].copy()


# ---- colour LUT ----------------------------------------------------------------

# BuiltinLUT gradients, aflak_plot/src/imshow/lut.rs:61-113
GRADIENTS = {
    "Grey": [(0.0, (0, 0, 0)), (1.0, (255, 255, 255))],
    "GreyClip": [(0.0, (0, 0, 0)), (0.99, (255, 255, 255)), (1.0, (255, 0, 0))],
    "Yellowy": [(0.0, (0, 0, 0)), (0.25, (32, 0, 129)), (0.5, (115, 15, 255)), (0.75, (255, 255, 0)), (1.0, (255, 255, 255))],
    "Thermal": [(0.0, (0, 0, 0)), (1.0 / 3.0, (185, 0, 0)), (2.0 / 3.0, (255, 220, 0)), (1.0, (255, 255, 255))],
    "Flame": [(0.0, (0, 0, 0)), (0.2, (7, 0, 220)), (0.5, (236, 0, 134)), (0.8, (246, 246, 0)), (1.0, (255, 255, 255))],
    "HeatMap": [(0.0, (1, 1, 85)), (0.1, (0, 0, 255)), (0.25, (0, 255, 255)), (0.5, (0, 255, 0)), (0.75, (255, 255, 0)),
                (0.9, (255, 0, 0)), (0.99, (108, 6, 10)), (1.0, (255, 255, 255))],
    "HeatMap_Inv": [(0.0, (108, 6, 10)), (0.1, (255, 0, 0)), (0.25, (255, 255, 0)), (0.5, (0, 255, 0)), (0.75, (0, 255, 255)),
                    (0.89, (0, 0, 255)), (0.99, (1, 1, 85)), (1.0, (255, 255, 255))],
    "Red": [(0.0, (0, 0, 0)), (1.0, (255, 0, 0))],
    "Green": [(0.0, (0, 0, 0)), (1.0, (0, 255, 0))],
    "Blue": [(0.0, (0, 0, 0)), (1.0, (0, 0, 255))],
}


def _stops(gradient):
    if isinstance(gradient, str):
        gradient = GRADIENTS[gradient]
    v = np.array([np.float32(s[0]) for s in gradient], dtype=np.float32)
    c = np.array([s[1] for s in gradient], dtype=np.uint8).reshape(-1, 3)
    return v, np.ascontiguousarray(c)


def _lims(lims):
    return (C.c_float * 3)(*[np.float32(x) for x in lims])


def lut_build(gradient, lims=(0.0, 0.5, 1.0)):
    v, c = _stops(gradient)
    lut = np.empty((LUT_SIZE, 3), dtype=np.uint8)
    lib().orc_lut_build(_fp(v), _u8p(c), C.c_int64(v.size), _lims(lims), _u8p(lut))
    return lut


def lut_bounds(gradient, lims=(0.0, 0.5, 1.0)):
    v, c = _stops(gradient)
    cap = v.size + 4
    v1 = np.empty(cap, np.float32)
    v2 = np.empty(cap, np.float32)
    c1 = np.empty((cap, 3), np.uint8)
    c2 = np.empty((cap, 3), np.uint8)
    n = lib().orc_lut_bounds(_fp(v), _u8p(c), C.c_int64(v.size), _lims(lims), _fp(v1), _u8p(c1), _fp(v2), _u8p(c2), C.c_int64(cap))
    return [((float(v1[i]), tuple(int(x) for x in c1[i])), (float(v2[i]), tuple(int(x) for x in c2[i]))) for i in range(n)]


def mtf(m, x):
    return np.float32(lib().orc_mtf(C.c_float(m), C.c_float(x)))


def lut_index(point, lims=(0.0, 0.5, 1.0)):
    return int(lib().orc_lut_index(C.c_float(point), _lims(lims)))


def color_at(lut, point, lims=(0.0, 0.5, 1.0)):
    rgb = (C.c_uint8 * 3)()
    lib().orc_color_at(_u8p(lut), _lims(lims), C.c_float(point), rgb)
    return [int(rgb[0]), int(rgb[1]), int(rgb[2])]


def color_at_bounds(lut, point, vmin, vmax, lims=(0.0, 0.5, 1.0)):
    rgb = (C.c_uint8 * 3)()
    lib().orc_color_at_bounds(_u8p(lut), _lims(lims), C.c_float(point), C.c_float(vmin), C.c_float(vmax), rgb)
    return [int(rgb[0]), int(rgb[1]), int(rgb[2])]


def get_vmin(img):
    img = _f32(img)
    v = C.c_float()
    if lib().orc_get_vmin(_fp(img), C.c_int64(img.size), C.byref(v)):
        raise OracleError(ERR_UNEXPECTED_INPUT, "Could not get vmin")
    return np.float32(v.value)


def get_vmax(img):
    img = _f32(img)
    v = C.c_float()
    if lib().orc_get_vmax(_fp(img), C.c_int64(img.size), C.byref(v)):
        raise OracleError(ERR_UNEXPECTED_INPUT, "Could not get vmax")
    return np.float32(v.value)


def make_raw_image(img, vmin, vmax, lut, lims=(0.0, 0.5, 1.0), channels=3, threads=0):
    img = _f32(img)
    ny, nx = img.shape
    out = np.empty((ny, nx, channels), dtype=np.uint8)
    if threads:
        _check(lib().orc_make_raw_image_mt(_fp(img), C.c_int64(ny), C.c_int64(nx), C.c_float(vmin), C.c_float(vmax), _u8p(lut), _lims(lims), C.c_int(channels), _u8p(out), C.c_int(threads)))
    else:
        _check(lib().orc_make_raw_image(_fp(img), C.c_int64(ny), C.c_int64(nx), C.c_float(vmin), C.c_float(vmax), _u8p(lut), _lims(lims), C.c_int(channels), _u8p(out)))
    return out


def make_raw_image_rgb(img, vmin3, vmax3, lut3, lims3, channels=3):
    img = _f32(img)
    _, ny, nx = img.shape
    lut3 = np.ascontiguousarray(lut3, dtype=np.uint8)
    lims3 = _f32(lims3).reshape(3, 3)
    vmin3, vmax3 = _f32(vmin3), _f32(vmax3)
    out = np.empty((ny, nx, channels), dtype=np.uint8)
    _check(lib().orc_make_raw_image_rgb(_fp(img), C.c_int64(ny), C.c_int64(nx), _fp(vmin3), _fp(vmax3), _u8p(lut3), _fp(lims3), C.c_int(channels), _u8p(out)))
    return out


# ---- synthetic generator --------------------------------------------------------

def gen_cube(seed, nz, ny, nx, y0=0, rows=None, pathology=False, threads=0):
    rows = ny - y0 if rows is None else rows
    out = np.empty((nz, rows, nx), dtype=np.float32)
    lib().orc_gen_cube(C.c_uint64(seed), C.c_int64(nz), C.c_int64(ny), C.c_int64(nx), C.c_int64(y0), C.c_int64(rows),
                       C.c_int(int(pathology)), _fp(out), C.c_int(threads or max_threads()))
    return out


# ---- display statistics (aflak_plot/src/imshow/hist.rs, lims.rs:37-106, zscale.rs) ---------------------------

def histogram(img, vmin, vmax, color=False):
    """hist::histogram (100 bins) / histogram_color ([3,ny,nx], 256 bins): (counts, starts, ends)."""
    img = _f32(img)
    nb, ch = (256, 3) if color else (100, 1)
    counts = np.zeros(nb * ch, dtype=np.uint64)
    lib().orc_histogram(_fp(img), C.c_int64(img.size // ch), C.c_int(ch), C.c_float(vmin), C.c_float(vmax), C.c_int(nb),
                        counts.ctypes.data_as(C.POINTER(C.c_uint64)))
    starts, ends = np.empty(nb, dtype=np.float32), np.empty(nb, dtype=np.float32)
    lib().orc_histogram_bins(C.c_float(vmin), C.c_float(vmax), C.c_int(nb), _fp(starts), _fp(ends))
    return (counts.reshape(nb, 3) if color else counts), starts, ends


def vmed_vmad(img):
    """(vmed, vmad) of lims::get_vmed_normalized / get_vmad_normalized; raises OracleError for the "Empty data!" cases."""
    img = _f32(img)
    vmin, vmax = get_vmin(img), get_vmax(img)
    med, mad = C.c_float(), C.c_float()
    if lib().orc_vmed_normalized(_fp(img), C.c_int64(img.size), C.c_float(vmin), C.c_float(vmax), C.byref(med)):
        raise OracleError(ERR_UNEXPECTED_INPUT, "Empty data!, med")
    if lib().orc_vmad_normalized(_fp(img), C.c_int64(img.size), C.c_float(vmin), C.c_float(vmax), med, C.byref(mad)):
        raise OracleError(ERR_UNEXPECTED_INPUT, "Empty data!, mad")
    return np.float32(med.value), np.float32(mad.value)


def zscale(img, nsamples=1000, contrast=0.25):
    img = _f32(img)
    assert img.ndim == 2
    has, z1, z2 = C.c_int(), C.c_float(), C.c_float()
    _check(lib().orc_zscale(_fp(img), C.c_int64(img.shape[0]), C.c_int64(img.shape[1]), C.c_int(nsamples), C.c_float(contrast),
                            C.byref(has), C.byref(z1), C.byref(z2)))
    return (np.float32(z1.value), np.float32(z2.value)) if has.value else (None, None)


# ---- spatial reductions / gathers and the remaining elementwise operators ----------------------------------------

def _roi(pixels):
    if pixels is None:
        return 1, None, 0, None
    a = np.ascontiguousarray(pixels, dtype=np.uint64).reshape(-1)
    return 0, a.ctypes.data_as(C.POINTER(C.c_uint64)), a.size // 2, a


def extract_wave(cube, pixels=None, f64=False):
    """run_extract_wave with ROI::All (pixels=None) or ROI::PixelList([(x, y), ...])."""
    cube = _f32(cube)
    nz, ny, nx = cube.shape
    roi_all, p, n, keep = _roi(pixels)
    out = np.empty(nz, dtype=np.float32)
    out64 = np.empty(nz, dtype=np.float64)
    _check(lib().orc_extract_wave(_fp(cube), C.c_int64(nz), C.c_int64(ny), C.c_int64(nx), C.c_int(roi_all), p, C.c_int64(n), _fp(out), _dp(out64)))
    return out64 if f64 else out


def extrude(cube, pixels=None):
    cube = _f32(cube)
    nz, ny, nx = cube.shape
    roi_all, p, n, keep = _roi(pixels)
    out = np.empty((nz, n), dtype=np.float32)
    _check(lib().orc_extrude(_fp(cube), C.c_int64(nz), C.c_int64(ny), C.c_int64(nx), C.c_int(roi_all), p, C.c_int64(n), _fp(out)))
    return out


def make_plane3d(p0, d1, d2, count1, count2):
    a, b, c = (np.ascontiguousarray(v, dtype=np.float32) for v in (p0, d1, d2))
    out = np.empty((count2, count1, 3), dtype=np.float32)
    lib().orc_make_plane3d(_fp(a), _fp(b), _fp(c), C.c_int64(count1), C.c_int64(count2), _fp(out))
    return out


def slice_3d_to_2d(cube, xyz):
    cube, xyz = _f32(cube), _f32(xyz)
    out = np.empty(xyz.shape[:2], dtype=np.float32)
    _check(lib().orc_slice_3d_to_2d(_fp(cube), C.c_int64(cube.shape[0]), C.c_int64(cube.shape[1]), C.c_int64(cube.shape[2]), _fp(xyz),
                                    C.c_int64(out.size), _fp(out)))
    return out


def velocity_field(img, w0):
    img = _f32(img)
    out = np.empty_like(img)
    lib().orc_velocity_field(_fp(img), C.c_int64(img.size), C.c_float(w0), _fp(out))
    return out


def histogram_transformation(img, parameters):
    img = _f32(img)
    p = np.ascontiguousarray(parameters, dtype=np.float32)
    out = np.empty_like(img)
    lib().orc_histogram_transformation(_fp(img), C.c_int64(img.size), _fp(p), _fp(out))
    return out


def histogram_transformation_rgb(img, parameters):
    img = _f32(img)
    p = np.ascontiguousarray(parameters, dtype=np.float32).reshape(-1, 3)
    return np.stack([histogram_transformation(img[c], p[c]) for c in range(img.shape[0])]) if img.shape[0] else img.copy()


def arcsinh_stretch(img, beta):
    img = _f32(img)
    out = img.copy()
    lib().orc_arcsinh_stretch(_fp(img), C.c_int64(img.shape[1] * img.shape[2]), C.c_float(beta), _fp(out))
    return out


def color_image_to_hsv(img):
    img = _f32(img)
    out = img.copy()
    lib().orc_color_image_to_hsv(_fp(img), C.c_int64(img.shape[1] * img.shape[2]), _fp(out))
    return out
