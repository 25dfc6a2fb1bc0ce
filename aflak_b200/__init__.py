"""aflak_b200 -- thin ctypes driver over libaflak_b200.so (the C ABI in include/aflak_b200.h).

The product is the shared library: hand-written sm_100a kernels behind a C ABI that mirrors aflak's
`cake::Transform` / `IOValue` operator registry.  This module only exists so that tests, bench.py and
examples can call that ABI from Python; it contains no compute and no fallback.  If the library is missing
the import builds it with nvcc (aflak_b200/build.py) and fails loudly when that is impossible.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libaflak_b200.so")
HEADER_PATH = os.path.join(os.path.dirname(_HERE), "include", "aflak_b200.h")

# Build on first import (or when the sources changed): nvcc for sm_100a, in-tree.  A failed build is a failed import --
# there is no other implementation to fall back to.
from . import build as _build  # noqa: E402

try:
    _build.build()
except Exception as exc:  # nvcc missing or a compile error
    raise ImportError(
        f"{LIB_PATH} could not be built ({exc}). The CUDA extension is the product; there is no CPU fallback."
    ) from exc
if not os.path.exists(LIB_PATH):
    raise ImportError(f"{LIB_PATH} not found: the CUDA extension is not built. There is no CPU fallback.")

lib = C.CDLL(LIB_PATH)

# ---- status / type enums (include/aflak_b200.h) -------------------------------------------------------
OK, ERR_IO, ERR_RAWLOADER, ERR_FITS, ERR_UNEXPECTED_INPUT, ERR_SHAPE, ERR_CUDA, ERR_PANIC, ERR_ARGUMENT = range(9)
(T_INTEGER, T_FLOAT, T_FLOAT2, T_FLOAT3, T_FLOAT3X3, T_STR, T_BOOL, T_FITS, T_IMAGE, T_MAP2DTO3DCOORDS, T_ROI, T_PATHS,
 T_TONECURVE) = range(13)
SUM_COMPENSATED, SUM_SEQUENTIAL = 0, 1
SUM_INHERIT = -1
EPI_SUM, EPI_MEAN, EPI_MEAN_RATIO, EPI_EW = 0, 1, 2, 3


class afk_result(C.Structure):
    _fields_ = [("status", C.c_int), ("value", C.c_void_p), ("message", C.c_void_p)]


class afk_input_slot(C.Structure):
    _fields_ = [("name", C.c_char_p), ("type_id", C.c_int), ("default_value", C.c_void_p)]


PLAIN_FUNCTION = C.CFUNCTYPE(C.c_int, C.POINTER(C.c_void_p), C.c_size_t, C.POINTER(afk_result), C.c_size_t)


class afk_transform_desc(C.Structure):
    _fields_ = [
        ("name", C.c_char_p), ("kind", C.c_char_p), ("description", C.c_char_p),
        ("major", C.c_uint8), ("minor", C.c_uint8), ("patch", C.c_uint8),
        ("n_inputs", C.c_size_t), ("inputs", C.POINTER(afk_input_slot)),
        ("n_outputs", C.c_size_t), ("outputs", C.POINTER(C.c_int)),
        ("f", PLAIN_FUNCTION),
    ]


class afk_colormap_chain(C.Structure):
    _fields_ = [("logscale_a", C.c_float), ("logscale_vmin", C.c_float), ("logscale_vmax", C.c_float),
                ("normalize_min", C.c_float), ("normalize_max", C.c_float),
                ("tone_table", C.POINTER(C.c_float)), ("tone_len", C.c_size_t)]


def _sig(name, restype, *argtypes):
    fn = getattr(lib, name)
    fn.restype = restype
    fn.argtypes = list(argtypes)
    return fn


_P = C.c_void_p
_i64p = C.POINTER(C.c_int64)
_fp = C.POINTER(C.c_float)
_u8p = C.POINTER(C.c_uint8)

_sig("afk_last_error", C.c_char_p)
_sig("afk_abi_version", C.c_int)
_sig("afk_device_count", C.c_int)
_sig("afk_set_device", C.c_int, C.c_int)
_sig("afk_set_stream", C.c_int, _P)
_sig("afk_get_stream", _P)
_sig("afk_synchronize", C.c_int)
_sig("afk_launch_count", C.c_uint64)
_sig("afk_type_name", C.c_char_p, C.c_int)
_sig("afk_type_from_name", C.c_int, C.c_char_p)
_sig("afk_type_convertible", C.c_int, C.c_int, C.c_int)
_sig("afk_type_editable", C.c_int, C.c_int)
_sig("afk_image_upload", _P, _P, C.c_int, _i64p)
_sig("afk_image_alloc", _P, C.c_int, _i64p)
_sig("afk_image_wrap_device", _P, _P, C.c_int, _i64p)
_sig("afk_image_synthetic", _P, C.c_uint64, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_int)
_sig("afk_image_retain", None, _P)
_sig("afk_image_release", None, _P)
_sig("afk_image_ndim", C.c_int, _P)
_sig("afk_image_shape", _i64p, _P)
_sig("afk_image_len", C.c_int64, _P)
_sig("afk_image_device_ptr", _P, _P)
_sig("afk_image_device", C.c_int, _P)
_sig("afk_image_download", C.c_int, _P, _P)
_sig("afk_image_unit", C.c_char_p, _P)
_sig("afk_image_set_unit", C.c_int, _P, C.c_char_p)
_sig("afk_image_homogeneous", C.c_int, _P)
_sig("afk_image_tag", C.c_char_p, _P)
_sig("afk_image_set_tag", C.c_int, _P, C.c_char_p)
_sig("afk_image_has_meta", C.c_int, _P)
_sig("afk_image_set_axis", C.c_int, _P, C.c_int, C.c_char_p, C.c_char_p)
_sig("afk_image_set_wcs_axis", C.c_int, _P, C.c_int, C.c_double, C.c_double, C.c_double)
_sig("afk_image_pix2world", C.c_int, _P, C.c_int, C.c_float, C.POINTER(C.c_float))
_sig("afk_image_axis_name", C.c_char_p, _P, C.c_int)
_sig("afk_image_axis_unit", C.c_char_p, _P, C.c_int)
_sig("afk_value_integer", _P, C.c_int64)
_sig("afk_value_float", _P, C.c_float)
_sig("afk_value_float2", _P, _fp)
_sig("afk_value_float3", _P, _fp)
_sig("afk_value_float3x3", _P, _fp)
_sig("afk_value_str", _P, C.c_char_p)
_sig("afk_value_bool", _P, C.c_int)
_sig("afk_value_image", _P, _P)
_sig("afk_value_tone_curve", _P, _fp, C.c_size_t, _fp, C.c_size_t)
_sig("afk_value_roi_all", _P)
_sig("afk_value_roi_pixels", _P, C.POINTER(C.c_uint64), C.c_size_t)
_sig("afk_value_roi_is_all", C.c_int, _P)
_sig("afk_value_roi_count", C.c_size_t, _P)
_sig("afk_value_roi_data", C.POINTER(C.c_uint64), _P)
_sig("afk_value_map2dto3d", _P, _fp, C.c_int64, C.c_int64)
_sig("afk_value_get_map", _fp, _P, _i64p, _i64p)
_sig("afk_value_fits", _P, _P)
_sig("afk_value_get_fits", _P, _P)
_sig("afk_value_paths", _P, C.POINTER(C.c_char_p), C.c_size_t)
_sig("afk_value_paths_count", C.c_size_t, _P)
_sig("afk_value_paths_at", C.c_char_p, _P, C.c_size_t)
_sig("afk_image_ingest", _P, _P, C.c_int, C.c_int, _i64p, C.c_int, C.c_int64, C.c_float, C.c_float)
_sig("afk_image_ingest_band", _P, _P, C.c_int, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_int, C.c_int64, C.c_float, C.c_float)
_sig("afk_fits_open", _P, C.c_char_p)
_sig("afk_fits_retain", None, _P)
_sig("afk_fits_release", None, _P)
_sig("afk_fits_hdu_count", C.c_size_t, _P)
_sig("afk_fits_header", C.c_char_p, _P, C.c_size_t, C.c_char_p)
_sig("afk_fits_to_image", C.c_int, _P, C.c_int64, C.c_char_p, C.POINTER(C.c_void_p))
_sig("afk_fits_to_image_band", C.c_int, _P, C.c_int64, C.c_char_p, C.c_int, C.c_int, C.POINTER(C.c_void_p))
_u64p = C.POINTER(C.c_uint64)
_sig("afk_image_histogram", C.c_int, _P, C.c_float, C.c_float, _u64p, _fp, _fp)
_sig("afk_image_histogram_color", C.c_int, _P, C.c_float, C.c_float, _u64p, _fp, _fp)
_sig("afk_image_vmed_vmad", C.c_int, _P, C.POINTER(C.c_float), C.POINTER(C.c_float))
_sig("afk_image_zscale", C.c_int, _P, C.c_int, C.c_float, C.POINTER(C.c_int), C.POINTER(C.c_float), C.POINTER(C.c_float))
_sig("afk_k_ingest", C.c_int, _P, C.c_int, C.c_int64, C.c_int, C.c_int64, C.c_float, C.c_float, _P)
_sig("afk_ingest_last_stats", C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_uint64))
_sig("afk_value_default_for", _P, C.c_int)
_sig("afk_value_retain", None, _P)
_sig("afk_value_release", None, _P)
_sig("afk_value_type", C.c_int, _P)
_sig("afk_value_get_integer", C.c_int, _P, _i64p)
_sig("afk_value_get_float", C.c_int, _P, _fp)
_sig("afk_value_get_floats", C.c_int, _P, _fp, C.c_size_t)
_sig("afk_value_get_bool", C.c_int, _P, C.POINTER(C.c_int))
_sig("afk_value_get_str", C.c_char_p, _P)
_sig("afk_value_get_image", _P, _P)
_sig("afk_value_convert", _P, _P, C.c_int)
_sig("afk_value_eq", C.c_int, _P, _P)
_sig("afk_result_clear", None, C.POINTER(afk_result))
_sig("afk_result_display", _P, C.POINTER(afk_result))
_sig("afk_free", None, _P)
_sig("afk_transform_count", C.c_size_t)
_sig("afk_transform_at", C.POINTER(afk_transform_desc), C.c_size_t)
_sig("afk_get_transform", C.POINTER(afk_transform_desc), C.c_char_p)
_sig("afk_transform_call", C.c_int, C.POINTER(afk_transform_desc), C.POINTER(C.c_void_p), C.c_size_t, C.POINTER(afk_result), C.c_size_t)
_sig("afk_set_sum_mode", C.c_int, C.c_int)
_sig("afk_get_sum_mode", C.c_int)
_sig("afk_set_default_sum_mode", C.c_int, C.c_int)
_sig("afk_trim_pools", C.c_int)
_sig("afk_fused_logscale_minmax", C.c_int, _P, C.c_float, C.POINTER(C.c_void_p))
_sig("afk_fused_average_ratio", C.c_int, _P, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.POINTER(C.c_void_p))
_sig("afk_fused_equivalent_width", C.c_int, _P, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_int64,
     C.c_float, C.c_float, C.c_int, C.POINTER(C.c_void_p), _fp)
_sig("afk_lut_linear", _P, _fp, _u8p, C.c_size_t)
_sig("afk_lut_builtin", _P, C.c_char_p)
_sig("afk_lut_builtin_count", C.c_size_t)
_sig("afk_lut_builtin_name", C.c_char_p, C.c_size_t)
_sig("afk_lut_release", None, _P)
_sig("afk_lut_set_lims", None, _P, C.c_float, C.c_float, C.c_float)
_sig("afk_lut_lims", None, _P, _fp)
_sig("afk_lut_set_gradient", C.c_int, _P, _fp, _u8p, C.c_size_t)
_sig("afk_lut_color_at", None, _P, C.c_float, _u8p)
_sig("afk_lut_color_at_bounds", None, _P, C.c_float, C.c_float, C.c_float, _u8p)
_sig("afk_lut_table", None, _P, _u8p)
_sig("afk_lut_bounds", C.c_size_t, _P, _fp, _u8p, _fp, _u8p, C.c_size_t)
_sig("afk_image_vmin_vmax", C.c_int, _P, _fp, _fp)
_sig("afk_set_display_exact", C.c_int, C.c_int)
_sig("afk_get_display_exact", C.c_int)
_sig("afk_colormap", C.c_int, _P, _P, C.c_float, C.c_float, C.c_int, C.c_int, C.POINTER(afk_colormap_chain), _P)
_sig("afk_colormap_host", C.c_int, _P, _P, C.c_float, C.c_float, C.c_int, C.c_int, C.POINTER(afk_colormap_chain), _P)
_sig("afk_colormap_rgb_host", C.c_int, _P, C.POINTER(C.c_void_p), _fp, _fp, C.c_int, C.c_int, _P)
_sig("afk_k_reduce", C.c_int, _P, C.c_int64, C.c_int64, C.c_int, _i64p, _i64p, C.c_int, _fp, C.c_int, _P)
_sig("afk_k_reduce_static", C.c_int, _P, C.c_int64, C.c_int64, C.c_int, _i64p, _i64p, C.c_int, _fp, C.c_int, _P, _P)
_sig("afk_k_reduce_describe", C.c_int, _P, C.c_int64, C.c_int, _i64p, _i64p, C.c_int, _P, C.c_char_p, C.c_size_t)
_sig("afk_profile_reduce", C.c_int, C.c_int)
_sig("afk_profile_reduce_read", C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_uint64))
_sig("afk_ipc_export", C.c_int, _P, _u8p)
_sig("afk_ipc_open", _P, _u8p)
_sig("afk_ipc_close", C.c_int, _P)
_sig("afk_shard_band", C.c_int, C.c_int64, C.c_int, C.c_int, _i64p, _i64p)
_sig("afk_tone_curve_from_points", C.c_int, _fp, C.c_size_t, _fp, C.c_size_t, C.POINTER(C.c_size_t))
_sig("afk_gather_create", _P, C.c_int, C.c_int, C.c_int64, C.c_int64, _u8p)
_sig("afk_gather_connect", C.c_int, _P, _u8p)
_sig("afk_gather_bytes", C.c_size_t, C.c_int64, C.c_int64)
_sig("afk_gather_create_symmetric", _P, C.c_int, C.c_int, C.c_int64, C.c_int64, C.POINTER(C.c_void_p), _P)
_sig("afk_gather_band", C.c_int, _P, C.c_int, _i64p, _i64p)
_sig("afk_gather_reduce", C.c_int, _P, _P, C.c_int64, C.c_int, _i64p, _i64p, C.c_int, _fp, C.c_int)
_sig("afk_gather_wait", C.c_int, _P, C.POINTER(C.c_void_p), C.c_double)
_sig("afk_gather_check", C.c_int, _P)
_sig("afk_gather_result_image", _P, _P)
_sig("afk_gather_colormap", C.c_int, _P, _P, C.c_float, C.c_float, C.c_int, _P)
_sig("afk_gather_wait_rgba", C.c_int, _P, C.POINTER(C.c_void_p), C.c_double)
_sig("afk_gather_destroy", None, _P)
_sig("afk_graph_new", _P)
_sig("afk_graph_free", None, _P)
_sig("afk_graph_add_transform", C.c_int64, _P, C.c_char_p)
_sig("afk_graph_add_constant", C.c_int64, _P, _P)
_sig("afk_graph_set_constant", C.c_int, _P, C.c_int64, _P)
_sig("afk_graph_connect", C.c_int, _P, C.c_int64, C.c_size_t, C.c_int64, C.c_size_t)
_sig("afk_graph_disconnect", C.c_int, _P, C.c_int64, C.c_size_t)
_sig("afk_graph_compute", C.c_int, _P, C.c_int64, C.c_size_t, C.POINTER(afk_result))
_sig("afk_graph_calls", C.c_uint64, _P)
_sig("afk_graph_set_fusion", None, _P, C.c_int)


class AflakError(Exception):
    """An `Err(IOErr)` (or a reference panic) coming back through the ABI."""

    def __init__(self, status, message):
        super().__init__(f"[{status}] {message}")
        self.status = status
        self.message = message


def _check(status):
    if status != OK:
        raise AflakError(status, lib.afk_last_error().decode())


def device_count():
    return lib.afk_device_count()


def set_device(i):
    _check(lib.afk_set_device(i))


def synchronize():
    _check(lib.afk_synchronize())


def set_stream(ptr):
    _check(lib.afk_set_stream(C.c_void_p(ptr) if ptr else None))


def launch_count():
    return int(lib.afk_launch_count())


def set_sum_mode(mode):
    """Per-thread override of the process-wide default; SUM_INHERIT (-1) removes it."""
    _check(lib.afk_set_sum_mode(mode))


def set_default_sum_mode(mode):
    _check(lib.afk_set_default_sum_mode(mode))


def get_sum_mode():
    return int(lib.afk_get_sum_mode())


def trim_pools():
    _check(lib.afk_trim_pools())


def type_name(t):
    return lib.afk_type_name(t).decode()


# ---- Image: device-resident WcsArray handle -------------------------------------------------------------
class Image:
    def __init__(self, handle, owned=True):
        if not handle:
            raise AflakError(ERR_CUDA, lib.afk_last_error().decode())
        self._h = C.c_void_p(handle)
        self._owned = owned

    @classmethod
    def upload(cls, array, sync=True):
        """sync=False: asynchronous copy from a pinned array the caller keeps alive until the stream has drained."""
        a = np.ascontiguousarray(array, dtype=np.float32)
        shape = (C.c_int64 * max(a.ndim, 1))(*a.shape)
        img = cls(lib.afk_image_upload(a.ctypes.data_as(_P), a.ndim, shape))
        if sync:
            synchronize()  # `a` may be a temporary
        return img

    @classmethod
    def synthetic(cls, seed, nz, ny, nx, y0=0, rows=None, pathology=False):
        rows = ny - y0 if rows is None else rows
        return cls(lib.afk_image_synthetic(seed, nz, ny, nx, y0, rows, int(pathology)))

    @classmethod
    def wrap_device(cls, ptr, shape):
        sh = (C.c_int64 * max(len(shape), 1))(*shape)
        return cls(lib.afk_image_wrap_device(C.c_void_p(ptr), len(shape), sh))

    @classmethod
    def alloc(cls, shape):
        sh = (C.c_int64 * max(len(shape), 1))(*shape)
        return cls(lib.afk_image_alloc(len(shape), sh))

    @property
    def handle(self):
        return self._h

    @property
    def shape(self):
        n = lib.afk_image_ndim(self._h)
        p = lib.afk_image_shape(self._h)
        return tuple(int(p[i]) for i in range(n))

    @property
    def device_ptr(self):
        return lib.afk_image_device_ptr(self._h)

    @property
    def unit(self):
        return lib.afk_image_unit(self._h).decode()

    @unit.setter
    def unit(self, u):
        lib.afk_image_set_unit(self._h, u.encode() if u else None)

    @property
    def tag(self):
        t = lib.afk_image_tag(self._h)
        return None if t is None else t.decode()

    @tag.setter
    def tag(self, t):
        lib.afk_image_set_tag(self._h, None if t is None else t.encode())

    @property
    def has_meta(self):
        return bool(lib.afk_image_has_meta(self._h))

    @property
    def homogeneous(self):
        return bool(lib.afk_image_homogeneous(self._h))

    def set_axis(self, i, name, unit):
        lib.afk_image_set_axis(self._h, i, None if name is None else name.encode(), (unit or "").encode())

    def set_wcs_axis(self, i, crval, crpix, cdelt):
        lib.afk_image_set_wcs_axis(self._h, i, crval, crpix, cdelt)

    def pix2world(self, axis, pixel):
        w = C.c_float()
        return w.value if lib.afk_image_pix2world(self._h, axis, pixel, C.byref(w)) else None

    def axis(self, i):
        return lib.afk_image_axis_name(self._h, i).decode(), lib.afk_image_axis_unit(self._h, i).decode()

    def download(self, out=None):
        if out is None:
            out = np.empty(self.shape, dtype=np.float32)
        assert out.dtype == np.float32 and out.flags.c_contiguous and out.size == int(np.prod(self.shape, dtype=np.int64))
        _check(lib.afk_image_download(self._h, out.ctypes.data_as(_P)))
        return out

    def histogram(self, vmin, vmax, color=False):
        """hist::histogram / histogram_color: (counts, starts, ends); counts is [100] or [256, 3]."""
        nb = 256 if color else 100
        counts = np.zeros(nb * (3 if color else 1), dtype=np.uint64)
        starts, ends = np.empty(nb, dtype=np.float32), np.empty(nb, dtype=np.float32)
        fn = lib.afk_image_histogram_color if color else lib.afk_image_histogram
        _check(fn(self._h, vmin, vmax, counts.ctypes.data_as(_u64p), starts.ctypes.data_as(_fp), ends.ctypes.data_as(_fp)))
        return (counts.reshape(nb, 3) if color else counts), starts, ends

    def vmed_vmad(self):
        a, b = C.c_float(), C.c_float()
        _check(lib.afk_image_vmed_vmad(self._h, C.byref(a), C.byref(b)))
        return np.float32(a.value), np.float32(b.value)

    def zscale(self, nsamples=1000, contrast=0.25):
        has, z1, z2 = C.c_int(), C.c_float(), C.c_float()
        _check(lib.afk_image_zscale(self._h, nsamples, contrast, C.byref(has), C.byref(z1), C.byref(z2)))
        return (np.float32(z1.value), np.float32(z2.value)) if has.value else (None, None)

    def vmin_vmax(self):
        a, b = C.c_float(), C.c_float()
        _check(lib.afk_image_vmin_vmax(self._h, C.byref(a), C.byref(b)))
        return np.float32(a.value), np.float32(b.value)

    def __del__(self):
        if getattr(self, "_owned", False) and self._h and lib is not None:  # lib is None during interpreter shutdown
            lib.afk_image_release(self._h)
            self._h = None


FITS_F32, FITS_F64, FITS_I16, FITS_I32, FITS_U32 = range(5)
_FITS_DTYPES = {FITS_F32: ">f4", FITS_F64: ">f8", FITS_I16: ">i2", FITS_I32: ">i4", FITS_U32: ">u4"}


def ingest(raw, fmt, shape, blank=None, bscale=1.0, bzero=0.0, band=None, sync=True):
    """Decode a big-endian FITS data unit held in host memory (afk_image_ingest / afk_image_ingest_band).

    raw: bytes-like or numpy array holding the data unit as it sits in the file; band = (y0, rows) of a 3-D cube."""
    buf = np.frombuffer(raw, dtype=np.uint8) if not isinstance(raw, np.ndarray) else raw.view(np.uint8).reshape(-1)
    has_blank, bl = (0, 0) if blank is None else (1, int(blank))
    if band is None:
        sh = (C.c_int64 * max(len(shape), 1))(*shape)
        h = lib.afk_image_ingest(buf.ctypes.data_as(_P), fmt, len(shape), sh, has_blank, bl, bscale, bzero)
    else:
        nz, ny, nx = shape
        h = lib.afk_image_ingest_band(buf.ctypes.data_as(_P), fmt, nz, ny, nx, band[0], band[1], has_blank, bl, bscale, bzero)
    img = Image(h)
    if sync:
        synchronize()  # `raw` may be a temporary
    return img


def k_ingest(dev_raw, fmt, n, dev_out, blank=None, bscale=1.0, bzero=0.0):
    _check(lib.afk_k_ingest(C.c_void_p(dev_raw), fmt, n, 0 if blank is None else 1, 0 if blank is None else int(blank), bscale, bzero,
                            C.c_void_p(dev_out)))


def ingest_stats():
    ms = (C.c_double * 3)()
    b = C.c_uint64()
    lib.afk_ingest_last_stats(ms, C.byref(b))
    return {"read_ms": ms[0], "wall_ms": ms[1], "device_ms": ms[2], "bytes": int(b.value)}


class Fits:
    """fitrs::Fits: parsed headers of every HDU; data units stay on disk until fits_to_image."""

    def __init__(self, handle, retain=False):
        if not handle:
            raise AflakError(ERR_IO, lib.afk_last_error().decode())
        self._h = C.c_void_p(handle)
        if retain:
            lib.afk_fits_retain(self._h)

    @classmethod
    def open(cls, path):
        return cls(lib.afk_fits_open(str(path).encode()))

    @property
    def handle(self):
        return self._h

    def __len__(self):
        return int(lib.afk_fits_hdu_count(self._h))

    def header(self, key, hdu=0):
        v = lib.afk_fits_header(self._h, hdu, key.encode())
        return None if v is None else v.decode()

    def to_image(self, hdu_idx=0, extension="", world=1, rank=0):
        out = C.c_void_p()
        _check(lib.afk_fits_to_image_band(self._h, hdu_idx, extension.encode(), world, rank, C.byref(out)))
        return Image(out.value)

    def __del__(self):
        if getattr(self, "_h", None) and lib is not None:
            lib.afk_fits_release(self._h)
            self._h = None


class Paths(list):
    """PATHS::FileList"""


class Roi:
    """roi::ROI: Roi() is ROI::All, Roi([(x, y), ...]) is ROI::PixelList (pairs are (i, j) = (column, row))."""

    def __init__(self, pixels=None):
        self.pixels = None if pixels is None else [(int(x), int(y)) for x, y in pixels]

    def __eq__(self, other):
        return isinstance(other, Roi) and self.pixels == other.pixels

    def __repr__(self):
        return "Roi(All)" if self.pixels is None else f"Roi({len(self.pixels)} pixels)"


class Map3d:
    """Map2dTo3dCoords: Array2<[f32; 3]> as a [rows, cols, 3] float32 array."""

    def __init__(self, xyz):
        self.xyz = np.ascontiguousarray(xyz, dtype=np.float32)
        assert self.xyz.ndim == 3 and self.xyz.shape[2] == 3


def tone_curve_from_points(cp):
    """UiToneCurve::create_curve (imgui_tone_curve/src/lib.rs:68-158): control points [(x, y), ...] -> table."""
    xy = np.ascontiguousarray(cp, dtype=np.float32).reshape(-1)
    n = C.c_size_t()
    _check(lib.afk_tone_curve_from_points(xy.ctypes.data_as(_fp), xy.size // 2, None, 0, C.byref(n)))
    out = np.empty(n.value, dtype=np.float32)
    _check(lib.afk_tone_curve_from_points(xy.ctypes.data_as(_fp), xy.size // 2, out.ctypes.data_as(_fp), out.size, C.byref(n)))
    return out


class ToneCurve:
    """ToneCurveState: `arr` table + control points (imgui_tone_curve/src/lib.rs:10-48)."""

    def __init__(self, arr=None, cp=((0.0, 0.0), (1.0, 1.0))):
        self.arr = np.arange(256, dtype=np.float32) if arr is None else np.ascontiguousarray(arr, dtype=np.float32)
        self.cp = np.ascontiguousarray(cp, dtype=np.float32).reshape(-1, 2)


class Value:
    """An IOValue converted once and kept (like a constant node or a cached output of the cake graph): passing it to
    several calls reuses the same afk_value, including whatever the library has attached to it (the device copy of a mesh)."""

    def __init__(self, x):
        self._v = C.c_void_p(_to_value(x))

    def __del__(self):
        if getattr(self, "_v", None) and lib is not None:
            lib.afk_value_release(self._v)
            self._v = None


def _to_value(x):
    """Python object -> new afk_value* (caller releases)."""
    if isinstance(x, Value):
        lib.afk_value_retain(x._v)
        return x._v.value
    if isinstance(x, Image):
        return lib.afk_value_image(x.handle)
    if isinstance(x, bool):
        return lib.afk_value_bool(int(x))
    if isinstance(x, (int, np.integer)):
        return lib.afk_value_integer(int(x))
    if isinstance(x, (float, np.floating)):
        return lib.afk_value_float(float(x))
    if isinstance(x, str):
        return lib.afk_value_str(x.encode())
    if isinstance(x, Roi):
        if x.pixels is None:
            return lib.afk_value_roi_all()
        a = np.ascontiguousarray(x.pixels, dtype=np.uint64).reshape(-1)
        return lib.afk_value_roi_pixels(a.ctypes.data_as(C.POINTER(C.c_uint64)), a.size // 2)
    if isinstance(x, Map3d):
        return lib.afk_value_map2dto3d(x.xyz.ctypes.data_as(_fp), x.xyz.shape[0], x.xyz.shape[1])
    if isinstance(x, Fits):
        return lib.afk_value_fits(x.handle)
    if isinstance(x, Paths):
        arr = (C.c_char_p * max(len(x), 1))(*[str(p).encode() for p in x])
        return lib.afk_value_paths(arr, len(x))
    if isinstance(x, ToneCurve):
        return lib.afk_value_tone_curve(x.arr.ctypes.data_as(_fp), x.arr.size, x.cp.ctypes.data_as(_fp), x.cp.shape[0])
    if isinstance(x, (tuple, list, np.ndarray)):
        a = np.ascontiguousarray(x, dtype=np.float32).reshape(-1)
        ctor = {2: lib.afk_value_float2, 3: lib.afk_value_float3, 9: lib.afk_value_float3x3}.get(a.size)
        if ctor is None:
            raise TypeError("float tuples must have 2, 3 or 9 entries")
        return ctor(a.ctypes.data_as(_fp))
    raise TypeError(f"cannot convert {type(x)} to an IOValue")


def _from_value(v):
    """afk_value* (borrowed) -> Python object. Images get their own reference."""
    t = lib.afk_value_type(v)
    if t == T_INTEGER:
        o = C.c_int64()
        lib.afk_value_get_integer(v, C.byref(o))
        return int(o.value)
    if t == T_FLOAT:
        o = C.c_float()
        lib.afk_value_get_float(v, C.byref(o))
        return np.float32(o.value)
    if t in (T_FLOAT2, T_FLOAT3, T_FLOAT3X3):
        buf = (C.c_float * 9)()
        n = lib.afk_value_get_floats(v, buf, 9)
        return np.array(buf[:n], dtype=np.float32)
    if t == T_BOOL:
        o = C.c_int()
        lib.afk_value_get_bool(v, C.byref(o))
        return bool(o.value)
    if t == T_STR:
        return lib.afk_value_get_str(v).decode()
    if t == T_IMAGE:
        h = lib.afk_value_get_image(v)
        lib.afk_image_retain(h)
        return Image(h)
    if t == T_TONECURVE:
        buf = (C.c_float * 4096)()
        n = lib.afk_value_get_floats(v, buf, 4096)
        return ToneCurve(np.array(buf[:n], dtype=np.float32))
    if t == T_ROI:
        if lib.afk_value_roi_is_all(v):
            return Roi()
        n = lib.afk_value_roi_count(v)
        d = lib.afk_value_roi_data(v)
        return Roi([(d[2 * i], d[2 * i + 1]) for i in range(n)])
    if t == T_MAP2DTO3DCOORDS:
        r, c = C.c_int64(), C.c_int64()
        p = lib.afk_value_get_map(v, C.byref(r), C.byref(c))
        n = r.value * c.value * 3
        return Map3d(np.ctypeslib.as_array(p, shape=(n,)).copy().reshape(r.value, c.value, 3) if n else np.zeros((r.value, c.value, 3), np.float32))
    if t == T_FITS:
        return Fits(lib.afk_value_get_fits(v), retain=True)
    if t == T_PATHS:
        return Paths(lib.afk_value_paths_at(v, i).decode() for i in range(lib.afk_value_paths_count(v)))
    return None


class Transform:
    """View of one registry entry (Algorithm::Function, aflak_cake/src/transform.rs:33-48)."""

    def __init__(self, desc_ptr):
        self._d = desc_ptr
        d = desc_ptr.contents
        self.name = d.name.decode()
        self.kind = d.kind.decode()
        self.description = d.description.decode()
        self.version = (d.major, d.minor, d.patch)
        self.inputs = []
        for i in range(d.n_inputs):
            s = d.inputs[i]
            default = _from_value(C.c_void_p(s.default_value)) if s.default_value else None
            self.inputs.append((s.name.decode(), type_name(s.type_id), default))
        self.outputs = [type_name(d.outputs[i]) for i in range(d.n_outputs)]

    def call_results(self, *args):
        """Feed + call; returns one entry per output: a Python value, or an AflakError instance."""
        d = self._d.contents
        vals = [None if a is None else _to_value(a) for a in args]
        arr = (C.c_void_p * max(len(vals), 1))(*[C.c_void_p(v) if v else None for v in vals])
        res = (afk_result * max(d.n_outputs, 1))()
        try:
            lib.afk_transform_call(self._d, arr, len(vals), res, d.n_outputs)
            out = []
            for i in range(d.n_outputs):
                r = res[i]
                if r.status == OK and r.value:
                    out.append(_from_value(C.c_void_p(r.value)))
                else:
                    msg = C.cast(r.message, C.c_char_p).value.decode() if r.message else lib.afk_last_error().decode()
                    out.append(AflakError(r.status if r.status != OK else ERR_PANIC, msg))
            return out
        finally:
            for i in range(d.n_outputs):
                lib.afk_result_clear(C.byref(res[i]))
            for v in vals:
                if v:
                    lib.afk_value_release(C.c_void_p(v))

    def __call__(self, *args):
        out = self.call_results(*args)
        for o in out:
            if isinstance(o, AflakError):
                raise o
        return out[0] if len(out) == 1 else tuple(out)


def get_transform(name):
    """NamedAlgorithms::get_transform (aflak_primitives/src/lib.rs:749-758)."""
    p = lib.afk_get_transform(name.encode())
    return Transform(p) if p else None


def transformations():
    return [Transform(lib.afk_transform_at(i)) for i in range(lib.afk_transform_count())]


def call(name, *args):
    t = get_transform(name)
    if t is None:
        raise KeyError(name)
    return t(*args)


# ---- fused paths --------------------------------------------------------------------------------------------
def fused_average_ratio(cube, s0, e0, s1, e1):
    out = C.c_void_p()
    _check(lib.afk_fused_average_ratio(cube.handle, s0, e0, s1, e1, C.byref(out)))
    return Image(out.value)


def fused_logscale_minmax(img, a):
    out = C.c_void_p()
    _check(lib.afk_fused_logscale_minmax(img.handle, C.c_float(a), C.byref(out)))
    return Image(out.value)


def fused_equivalent_width(cube, on, off1, off2, fl, vmax, is_emission, want_minmax=False):
    out = C.c_void_p()
    mm = (C.c_float * 2)()
    _check(lib.afk_fused_equivalent_width(cube.handle, on[0], on[1], off1[0], off1[1], off2[0], off2[1], fl, vmax,
                                          int(is_emission), C.byref(out), mm if want_minmax else None))
    img = Image(out.value)
    return (img, (np.float32(mm[0]), np.float32(mm[1]))) if want_minmax else img


def _reduce_params(epilogue, params):
    """EPI_EW takes exactly (a, b, fl, max, sign); the other epilogues read no params."""
    if epilogue != EPI_EW:
        return None
    if params is None or len(params) != 5:
        raise ValueError("the equivalent-width epilogue needs exactly 5 params: a, b, fl, max, sign")
    return (C.c_float * 5)(*params)


def k_reduce(cube_ptr, nz, n2, windows, epilogue, out_ptr, params=None, sum_mode=SUM_COMPENSATED):
    """Raw kernel layer: device pointers in, launch on this thread's stream."""
    n = len(windows)
    st = (C.c_int64 * n)(*[w[0] for w in windows])
    en = (C.c_int64 * n)(*[w[1] for w in windows])
    pr = _reduce_params(epilogue, params)
    _check(lib.afk_k_reduce(C.c_void_p(cube_ptr), nz, n2, n, st, en, epilogue, pr, sum_mode, C.c_void_p(out_ptr)))


def k_reduce_static(cube_ptr, nz, n2, windows, epilogue, out_ptr, params=None, sum_mode=SUM_COMPENSATED, stream=None):
    """k_reduce without any allocation: the split-K scratch is a static __device__ array of the library (afk_k_reduce_static).
    stream: a cudaStream_t handle (None = the legacy default stream)."""
    n = len(windows)
    st = (C.c_int64 * n)(*[w[0] for w in windows])
    en = (C.c_int64 * n)(*[w[1] for w in windows])
    pr = _reduce_params(epilogue, params)
    _check(lib.afk_k_reduce_static(C.c_void_p(cube_ptr), nz, n2, n, st, en, epilogue, pr, sum_mode, C.c_void_p(out_ptr),
                                   C.c_void_p(stream) if stream else None))


def k_reduce_describe(cube_ptr, n2, windows, out_ptr, sum_mode=SUM_COMPENSATED):
    """The kernel instantiation + launch shape k_reduce / Gather.reduce use for this problem, as text."""
    n = len(windows)
    st = (C.c_int64 * n)(*[w[0] for w in windows])
    en = (C.c_int64 * n)(*[w[1] for w in windows])
    buf = C.create_string_buffer(256)
    _check(lib.afk_k_reduce_describe(C.c_void_p(cube_ptr), n2, n, st, en, sum_mode, C.c_void_p(out_ptr), buf, 256))
    return buf.value.decode()


def profile_reduce(enable):
    lib.afk_profile_reduce(int(enable))


def profile_reduce_read():
    ms, n = C.c_double(), C.c_uint64()
    _check(lib.afk_profile_reduce_read(C.byref(ms), C.byref(n)))
    return ms.value, int(n.value)


# ---- spatial sharding + fused gather (one process per GPU) ----------------------------------------------------
def gather_bytes(ny_total, nx):
    return int(lib.afk_gather_bytes(ny_total, nx))


def symmetric_memory_torch(nbytes, group=None):
    """Host-side plumbing for Gather(symmetric=...): a zeroed symmetric allocation through torch's symmetric-memory allocator
    (cuMem + NVLink multicast object + handle exchange over the process group's store).  Returns (peer_ptrs, multicast_ptr,
    keep_alive); multicast_ptr is 0 where the fabric has no NVLink-SHARP multicast.  Collective: every rank calls it."""
    import torch
    import torch.distributed as dist
    import torch.distributed._symmetric_memory as symm_mem

    group = group or dist.group.WORLD
    t = symm_mem.empty(int(nbytes), dtype=torch.uint8, device=torch.device("cuda", torch.cuda.current_device()))
    t.zero_()
    h = symm_mem.rendezvous(t, group)
    torch.cuda.synchronize()
    dist.barrier(group)
    return [int(p) for p in h.buffer_ptrs], int(h.multicast_ptr), (t, h)


def shard_band(ny, world, rank):
    """y-band [y0, y0+rows) of `rank`: pure host arithmetic, usable without a GPU."""
    y0, rows = C.c_int64(), C.c_int64()
    _check(lib.afk_shard_band(ny, world, rank, C.byref(y0), C.byref(rows)))
    return int(y0.value), int(rows.value)


class Gather:
    """Fused gather of the 2-D result bands across the ranks of one box (include/aflak_b200.h, afk_gather_*).

    `exchange(handle_bytes) -> list of every rank's handle bytes` is supplied by the host side's process group
    (e.g. torch.distributed.all_gather_object); only 64 bytes per rank cross it, once.
    """

    def __init__(self, world, rank, ny_total, nx, exchange=None, symmetric=None):
        """symmetric = (peer_ptrs, multicast_ptr, keep_alive): host-supplied symmetric memory of gather_bytes(ny, nx) zeroed
        bytes per rank (afk_gather_create_symmetric); otherwise the library allocates and `exchange` carries the IPC handles."""
        self.world, self.rank, self.ny, self.nx = world, rank, ny_total, nx
        if symmetric is not None:
            peers, mc, self._keep = symmetric
            arr = (C.c_void_p * world)(*[C.c_void_p(int(p)) for p in peers])
            self._h = C.c_void_p(lib.afk_gather_create_symmetric(world, rank, ny_total, nx, arr, C.c_void_p(int(mc) if mc else None)))
            if not self._h:
                raise AflakError(ERR_CUDA, lib.afk_last_error().decode())
            self.multicast = bool(mc)
            return
        self.multicast = False
        h = (C.c_uint8 * 64)()
        self._h = C.c_void_p(lib.afk_gather_create(world, rank, ny_total, nx, h))
        if not self._h:
            raise AflakError(ERR_CUDA, lib.afk_last_error().decode())
        if world > 1:
            if exchange is None:
                raise ValueError("world > 1 needs an exchange function")
            allh = exchange(bytes(h))
            buf = (C.c_uint8 * (64 * world)).from_buffer_copy(b"".join(allh))
            _check(lib.afk_gather_connect(self._h, buf))

    def band(self, rank=None):
        y0, rows = C.c_int64(), C.c_int64()
        lib.afk_gather_band(self._h, self.rank if rank is None else rank, C.byref(y0), C.byref(rows))
        return int(y0.value), int(rows.value)

    def reduce(self, band_cube_ptr, nz, windows, epilogue, params=None, sum_mode=SUM_COMPENSATED):
        n = len(windows)
        st = (C.c_int64 * n)(*[w[0] for w in windows])
        en = (C.c_int64 * n)(*[w[1] for w in windows])
        pr = _reduce_params(epilogue, params)
        _check(lib.afk_gather_reduce(self._h, C.c_void_p(band_cube_ptr), nz, n, st, en, epilogue, pr, sum_mode))

    def wait(self, timeout_s=5.0):
        """Enqueue the wait on this thread's stream; returns the device pointer of the full [ny, nx] image."""
        p = C.c_void_p()
        _check(lib.afk_gather_wait(self._h, C.byref(p), timeout_s))
        return p.value

    def check(self):
        _check(lib.afk_gather_check(self._h))

    def result_image(self):
        return Image(lib.afk_gather_result_image(self._h))

    def colormap(self, lut, vmin=float("nan"), vmax=float("nan"), flip_y=True, chain=None):
        """Colour-map this rank's band of the gathered image and gather the RGBA8 bands on every rank (afk_gather_colormap)."""
        st, keep = _chain_struct(chain)
        _check(lib.afk_gather_colormap(self._h, lut.handle, vmin, vmax, int(flip_y), C.byref(st) if st else None))
        del keep

    def display(self, lut, vmin=float("nan"), vmax=float("nan"), flip_y=True, chain=None):
        """colormap + wait_rgba + download: this rank's full [ny, nx, 4] uint8 image on the host."""
        self.colormap(lut, vmin, vmax, flip_y, chain)
        ptr = self.wait_rgba()
        if self.ny * self.nx == 0:
            return np.empty((self.ny, self.nx, 4), dtype=np.uint8)
        # 4 bytes per pixel: read the byte image back through a borrowed f32 view of the same memory
        px = Image.wrap_device(ptr, (self.ny, self.nx)).download()
        return px.view(np.uint8).reshape(self.ny, self.nx, 4)

    def wait_rgba(self, timeout_s=5.0):
        """Enqueue the wait; returns the device pointer of this rank's full [ny, nx, 4] uint8 image."""
        p = C.c_void_p()
        _check(lib.afk_gather_wait_rgba(self._h, C.byref(p), timeout_s))
        return p.value

    def close(self):
        if getattr(self, "_h", None) and lib is not None:
            lib.afk_gather_destroy(self._h)
            self._h = None

    def __del__(self):
        self.close()


# ---- display path ---------------------------------------------------------------------------------------------
class Lut:
    """ColorLUT (aflak_plot/src/imshow/lut.rs)."""

    def __init__(self, gradient):
        if isinstance(gradient, str):
            self._h = C.c_void_p(lib.afk_lut_builtin(gradient.encode()))
        else:
            v = np.array([np.float32(s[0]) for s in gradient], dtype=np.float32)
            c = np.ascontiguousarray([s[1] for s in gradient], dtype=np.uint8).reshape(-1, 3)
            self._h = C.c_void_p(lib.afk_lut_linear(v.ctypes.data_as(_fp), c.ctypes.data_as(_u8p), v.size))
        if not self._h:
            raise AflakError(ERR_UNEXPECTED_INPUT, lib.afk_last_error().decode())

    @staticmethod
    def builtin_names():
        return [lib.afk_lut_builtin_name(i).decode() for i in range(lib.afk_lut_builtin_count())]

    @property
    def handle(self):
        return self._h

    def set_lims(self, mn, mid, mx):
        lib.afk_lut_set_lims(self._h, mn, mid, mx)

    def lims(self):
        out = (C.c_float * 3)()
        lib.afk_lut_lims(self._h, out)
        return tuple(np.float32(x) for x in out)

    def set_gradient(self, gradient):
        v = np.array([np.float32(s[0]) for s in gradient], dtype=np.float32)
        c = np.ascontiguousarray([s[1] for s in gradient], dtype=np.uint8).reshape(-1, 3)
        lib.afk_lut_set_gradient(self._h, v.ctypes.data_as(_fp), c.ctypes.data_as(_u8p), v.size)

    def color_at(self, point):
        rgb = (C.c_uint8 * 3)()
        lib.afk_lut_color_at(self._h, point, rgb)
        return [int(x) for x in rgb]

    def color_at_bounds(self, point, vmin, vmax):
        rgb = (C.c_uint8 * 3)()
        lib.afk_lut_color_at_bounds(self._h, point, vmin, vmax, rgb)
        return [int(x) for x in rgb]

    def table(self):
        out = np.empty((65536, 3), dtype=np.uint8)
        lib.afk_lut_table(self._h, out.ctypes.data_as(_u8p))
        return out

    def bounds(self):
        cap = 32
        v1 = (C.c_float * cap)()
        v2 = (C.c_float * cap)()
        c1 = (C.c_uint8 * (3 * cap))()
        c2 = (C.c_uint8 * (3 * cap))()
        n = lib.afk_lut_bounds(self._h, v1, c1, v2, c2, cap)
        return [((float(v1[i]), tuple(c1[3 * i:3 * i + 3])), (float(v2[i]), tuple(c2[3 * i:3 * i + 3]))) for i in range(n)]

    def __del__(self):
        if getattr(self, "_h", None) and lib is not None:
            lib.afk_lut_release(self._h)
            self._h = None


def _chain_struct(chain):
    if chain is None:
        return None, None
    table = np.ascontiguousarray(chain["tone_table"], dtype=np.float32)
    st = afk_colormap_chain(chain["logscale_a"], chain["logscale_vmin"], chain["logscale_vmax"], chain["normalize_min"],
                            chain["normalize_max"], table.ctypes.data_as(_fp), table.size)
    return st, table


def set_display_exact(on):
    """True: every pixel of the display kernels takes the reference's own IEEE sequence (the baseline of the parity tests)."""
    lib.afk_set_display_exact(int(bool(on)))


def colormap(img, lut, vmin=float("nan"), vmax=float("nan"), flip_y=True, channels=4, chain=None):
    """make_raw_image (aflak_plot/src/imshow/image.rs:20-49) -> host uint8 array [ny, nx, channels]."""
    ny, nx = img.shape
    out = np.empty((ny, nx, channels), dtype=np.uint8)
    st, keep = _chain_struct(chain)
    _check(lib.afk_colormap_host(img.handle, lut.handle, vmin, vmax, int(flip_y), channels, C.byref(st) if st else None,
                                 out.ctypes.data_as(_P)))
    del keep
    return out


def colormap_device(img, lut, vmin, vmax, out_ptr, flip_y=True, channels=4, chain=None):
    st, keep = _chain_struct(chain)
    _check(lib.afk_colormap(img.handle, lut.handle, vmin, vmax, int(flip_y), channels, C.byref(st) if st else None,
                            C.c_void_p(out_ptr)))
    del keep


def colormap_rgb(img, luts, vmin3, vmax3, flip_y=True, channels=3):
    _, ny, nx = img.shape
    out = np.empty((ny, nx, channels), dtype=np.uint8)
    hs = (C.c_void_p * 3)(*[l.handle for l in luts])
    a = (C.c_float * 3)(*vmin3)
    b = (C.c_float * 3)(*vmax3)
    _check(lib.afk_colormap_rgb_host(img.handle, hs, a, b, int(flip_y), channels, out.ctypes.data_as(_P)))
    return out


# ---- mini evaluator ---------------------------------------------------------------------------------------------
class Graph:
    """DST + Cache (aflak_cake/src/dst, cache.rs): device handles are what is cached."""

    def __init__(self):
        self._h = C.c_void_p(lib.afk_graph_new())

    def add_transform(self, name):
        i = lib.afk_graph_add_transform(self._h, name.encode())
        if i < 0:
            raise AflakError(ERR_UNEXPECTED_INPUT, lib.afk_last_error().decode())
        return int(i)

    def add_constant(self, x):
        v = _to_value(x)
        try:
            return int(lib.afk_graph_add_constant(self._h, C.c_void_p(v)))
        finally:
            lib.afk_value_release(C.c_void_p(v))

    def set_constant(self, t_idx, x):
        v = _to_value(x)
        try:
            _check(lib.afk_graph_set_constant(self._h, t_idx, C.c_void_p(v)))
        finally:
            lib.afk_value_release(C.c_void_p(v))

    def connect(self, from_t, out_i, to_t, in_i):
        _check(lib.afk_graph_connect(self._h, from_t, out_i, to_t, in_i))

    def disconnect(self, to_t, in_i):
        _check(lib.afk_graph_disconnect(self._h, to_t, in_i))

    def compute(self, t_idx, out_i=0):
        r = afk_result()
        st = lib.afk_graph_compute(self._h, t_idx, out_i, C.byref(r))
        try:
            if st != OK:
                msg = C.cast(r.message, C.c_char_p).value.decode() if r.message else lib.afk_last_error().decode()
                raise AflakError(st, msg)
            return _from_value(C.c_void_p(r.value))
        finally:
            lib.afk_result_clear(C.byref(r))

    @property
    def calls(self):
        return int(lib.afk_graph_calls(self._h))

    def set_fusion(self, enable):
        lib.afk_graph_set_fusion(self._h, int(enable))

    def __del__(self):
        if getattr(self, "_h", None) and lib is not None:
            lib.afk_graph_free(self._h)
            self._h = None
