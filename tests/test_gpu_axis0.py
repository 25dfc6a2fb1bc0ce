"""GPU parity: the other axis-0 reductions through the operator registry and the C ABI.

Reference: aflak_primitives/src/lib.rs:412-479 (registry), 1332-1582 (run_variance, run_stddev, run_median, run_minmax,
run_create_argmap, run_centroid, run_centroid_with_mask).

Tiers: median, arg maps, the max/min fold and (in SUM_SEQUENTIAL mode) variance / stddev / centroid are bit-exact against the
oracle's restatement of the reference's own f32 order.  In the default mode variance / stddev / centroid use f64 accumulators
and must be within 1e-5 relative of the f64 oracle (variance: two-pass f64; centroid: f64 quotient of the same products).
"""
import numpy as np
import pytest

from util import assert_bit_equal

pytestmark = pytest.mark.gpu

SHAPES = [
    (64, 8, 8),
    (200, 49, 49),     # odd n2 -> scalar lanes
    (131, 37, 52),     # n2 % 4 == 0, ragged tile
    (457, 74, 74),     # MaNGA spatial size: the spectral range is split between CTAs
    (33, 1, 1),
    (40, 300, 301),
]
WCS = (3621.6, 1.0, 0.83)  # CRVAL3, CRPIX3, CDELT3 of a MaNGA-like cube


def ranges(nz):
    out = [(0, nz), (nz // 3, 2 * nz // 3 + 1), (1, min(nz, 10)), (nz - 2, nz)]
    return [(a, b) for a, b in out if b - a >= 2]


def with_wave_axis(gpu, host, wcs=WCS):
    cube = gpu.Image.upload(host)
    cube.unit = "1E-17 erg/s/cm^2/Ang"
    cube.set_axis(0, "RA---TAN", "deg")
    cube.set_axis(1, "DEC--TAN", "deg")
    cube.set_axis(2, "WAVE", "Angstrom")
    if wcs is not None:
        cube.set_wcs_axis(2, *wcs)
    return cube


def assert_rel(got, want64, rtol, what):
    got = np.asarray(got, dtype=np.float64)
    assert np.array_equal(np.isnan(got), np.isnan(want64)), f"{what}: NaN positions differ"
    fin = np.isfinite(want64)
    assert np.array_equal(got[~fin & ~np.isnan(want64)], want64[~fin & ~np.isnan(want64)]), f"{what}: Inf positions differ"
    err = np.abs(got[fin] - want64[fin])
    tol = rtol * np.abs(want64[fin]) + 1e-37
    assert np.all(err <= tol), f"{what}: worst rel err {np.max(err / np.maximum(np.abs(want64[fin]), 1e-300))}"


@pytest.mark.parametrize("shape", SHAPES)
@pytest.mark.parametrize("name", ["variance", "stddev"])
def test_variance_default_mode_vs_f64(gpu, orc, shape, name):
    nz, ny, nx = shape
    host = orc.gen_cube(0xAF1A0011, nz, ny, nx)
    cube = gpu.Image.upload(host)
    for start, end in ranges(nz):
        got = gpu.call(name, cube, start, end).download()
        want = orc.variance_f64(host, start, end)
        if name == "stddev":
            want = np.sqrt(want)
        assert got.shape == (ny, nx)
        assert_rel(got, want, 1e-5, f"{name}{shape}[{start},{end})")


@pytest.mark.parametrize("shape", SHAPES)
@pytest.mark.parametrize("name", ["variance", "stddev"])
def test_variance_sequential_bit_exact(gpu, orc, shape, name):
    nz, ny, nx = shape
    host = orc.gen_cube(0xAF1A0012, nz, ny, nx)
    cube = gpu.Image.upload(host)
    gpu.set_sum_mode(gpu.SUM_SEQUENTIAL)
    try:
        for start, end in ranges(nz):
            got = gpu.call(name, cube, start, end).download()
            want = getattr(orc, name)(host, start, end)
            assert_bit_equal(got, want, f"{name}{shape}[{start},{end})")
    finally:
        gpu.set_sum_mode(gpu.SUM_COMPENSATED)


def test_variance_errors_and_meta(gpu, orc):
    host = orc.gen_cube(0xAF1A0013, 30, 6, 8)
    cube = with_wave_axis(gpu, host)
    # ndarray panics when ddof (1.0) >= frame count
    e = gpu.get_transform("variance").call_results(cube, 3, 4)[0]
    assert isinstance(e, gpu.AflakError) and e.status == gpu.ERR_PANIC and "ddof" in str(e)
    for args, msg in [((-1, 4), "'start' must be positive, but got -1"), ((5, 3), "Expected start < end, but got 5 < 3"),
                      ((0, 31), "'end' greater or equal to the input image 'image''s frame count (expected 31 < 31)")]:
        for name in ["variance", "stddev", "median", "extract_centrobaric_wavelength"]:
            e = gpu.get_transform(name).call_results(cube, *args)[0]
            assert isinstance(e, gpu.AflakError) and e.status == gpu.ERR_UNEXPECTED_INPUT and msg in str(e), (name, args, str(e))
    # reduce_array_slice -> make_slice: unit kept, first two axes kept, tag dropped
    v = gpu.call("variance", cube, 0, 30)
    assert v.unit == "1E-17 erg/s/cm^2/Ang" and v.has_meta and v.axis(0) == ("RA---TAN", "deg") and v.axis(2) == ("", "")
    # median: from_array(.., Unit::None)
    m = gpu.call("median", cube, 0, 30)
    assert m.unit == "" and not m.has_meta


@pytest.mark.parametrize("shape", SHAPES)
def test_median_bit_exact(gpu, orc, shape):
    nz, ny, nx = shape
    host = orc.gen_cube(0xAF1A0014, nz, ny, nx)
    # duplicates, zeros of both signs, negatives and denormals (no NaN: the reference panics on those)
    host[::7] = np.float32(1.25)
    host[1::11] = -host[1::11]
    host[2::13] = 0.0
    host[3::17] = -0.0
    host[4::19] = 1e-40
    cube = gpu.Image.upload(host)
    # windows <= 128 frames go through the sorting-network kernel (N = 8..128), longer ones through the radix select
    for start, end in [(0, nz), (0, nz - 1), (nz // 3, 2 * nz // 3 + 1), (5, 6), (1, 3), (0, min(nz, 64)), (0, min(nz, 65)), (2, min(nz, 33)),
                       (0, min(nz, 8)), (1, min(nz, 10)), (0, min(nz, 16)), (0, min(nz, 17)), (0, min(nz, 128)), (0, min(nz, 129)), (3, min(nz, 130))]:
        got = gpu.call("median", cube, start, end).download()
        want = orc.median(host, start, end)
        assert got.shape == (ny, nx)
        assert_bit_equal(got, want, f"median{shape}[{start},{end})")


def test_median_long_range_and_nan(gpu, orc):
    # a range long enough that only narrow tiles fit in shared memory
    nz, ny, nx = 9000, 4, 8
    host = orc.gen_cube(0xAF1A0015, nz, ny, nx)
    cube = gpu.Image.upload(host)
    for start, end in [(0, nz), (1, nz)]:
        assert_bit_equal(gpu.call("median", cube, start, end).download(), orc.median(host, start, end), "median long")
    # a range beyond shared memory: the column is re-read from global memory on every pass
    nz, ny, nx = 60001, 1, 3
    host = orc.gen_cube(0xAF1A001C, nz, ny, nx)
    host[::5] = -host[::5]
    cube = gpu.Image.upload(host)
    for start, end in [(0, nz), (1, nz)]:
        assert_bit_equal(gpu.call("median", cube, start, end).download(), orc.median(host, start, end), "median global")
    # constant spectra (span 0) and two-valued spectra
    host = np.full((100, 3, 4), 2.5, np.float32)
    host[::2, 1] = 7.0
    assert_bit_equal(gpu.call("median", gpu.Image.upload(host), 0, 100).download(), orc.median(host, 0, 100), "median constant")
    assert_bit_equal(gpu.call("median", gpu.Image.upload(host), 0, 99).download(), orc.median(host, 0, 99), "median constant odd")
    # the +-Inf values of the pathology cube sort like any other value; NaN -> the reference panics
    host = orc.gen_cube(0xAF1A0016, 200, 10, 12, pathology=True)
    clean = np.where(np.isnan(host), np.float32(0.5), host)
    assert np.isinf(clean).any()
    assert_bit_equal(gpu.call("median", gpu.Image.upload(clean), 0, 200).download(), orc.median(clean, 0, 200), "median inf")
    e = gpu.get_transform("median").call_results(gpu.Image.upload(host), 0, 200)[0]
    assert isinstance(e, gpu.AflakError) and e.status == gpu.ERR_PANIC
    # a single frame is never compared: NaN passes through
    one = gpu.call("median", gpu.Image.upload(host), 7, 8).download()
    assert_bit_equal(one, host[7], "median of one frame")


def _median_patterns(rng, nz, ny, nx):
    """Spectra that exercise every branch of the selects: (tag, cube)."""
    shape = (nz, ny, nx)
    out = []
    # smooth noise: the median's bin holds a few dozen to a few hundred keys
    out.append(("normal", (1.0 + 0.1 * rng.standard_normal(shape)).astype(np.float32)))
    # heavy outliers stretch the span: equal-width digits put nearly every key into the median's bin
    h = (1.0 + 1e-3 * rng.standard_normal(shape)).astype(np.float32)
    h[::97] = 1e30
    h[5::101] = -1e30
    out.append(("outliers", h))
    # more duplicates of the median value than any candidate list holds; the successor lives in a higher bin
    h = np.full(shape, 3.0, np.float32)
    k0, k1 = int(0.42 * nz), int(0.9 * nz)
    h[:k0] = rng.uniform(0.0, 2.9, (k0, ny, nx)).astype(np.float32)
    h[k1:] = rng.uniform(3.1, 9.0, (nz - k1, ny, nx)).astype(np.float32)
    out.append(("duplicates", rng.permutation(h, axis=0)))
    # exactly half below / half above a gap: n even -> the two middles come from different bins
    h = np.concatenate([rng.uniform(1.0, 1.5, (nz // 2, ny, nx)), rng.uniform(100.0, 200.0, (nz - nz // 2, ny, nx))]).astype(np.float32)
    out.append(("gap", rng.permutation(h, axis=0)))
    # neighbouring floats only (span below one radix digit), many ties
    steps = rng.integers(0, 100, shape).astype(np.uint32)
    out.append(("tiny span", (np.full(shape, 1.0, np.float32).view(np.uint32) + steps).view(np.float32)))
    # signed values, zeros of both signs, a sorted ramp along the window (stratified samples are exact quantiles)
    h = rng.standard_normal(shape).astype(np.float32)
    h[::9] = 0.0
    h[4::9] = -0.0
    out.append(("signed zeros", h))
    out.append(("ramp", (np.arange(nz, dtype=np.float32)[:, None, None] * np.float32(0.25) + rng.uniform(0, 0.1, (1, ny, nx)).astype(np.float32))))
    return out


def test_median_select_paths(gpu, orc):
    """Shared-memory select (short windows / small images): candidate bins, ties, successor in another bin, ragged tiles."""
    rng = np.random.default_rng(0xAF1A0041)
    for nz, ny, nx in [(301, 16, 16), (1200, 8, 8), (130, 3, 12), (90, 20, 52)]:
        for tag, host in _median_patterns(rng, nz, ny, nx):
            cube = gpu.Image.upload(host)
            for start, end in [(0, nz), (0, nz - 1), (3, nz - 2)]:
                assert_bit_equal(gpu.call("median", cube, start, end).download(), orc.median(host, start, end), f"{tag} {nz}x{ny}x{nx}[{start},{end})")


def test_median_stream_paths(gpu, orc):
    """Streaming select (windows > 64 frames on images of >= 9472 spaxels): the same patterns; 97 x 99 spaxels is not a
    multiple of the 32-spaxel tile.  300 frames fit the 768-slot list whole (no pivots); 2100 frames take 7 pivots,
    `duplicates` / `tiny span` end on the equal-to-pivot branch; the 6001-frame window takes 15 pivots and some bins still
    overflow the list (refinement against the global column)."""
    rng = np.random.default_rng(0xAF1A0042)
    ny, nx = 97, 99
    for nz in (300, 2100):
        for tag, host in _median_patterns(rng, nz, ny, nx):
            cube = gpu.Image.upload(host)
            for start, end in [(0, nz), (1, nz), (10, nz - 40)]:
                assert_bit_equal(gpu.call("median", cube, start, end).download(), orc.median(host, start, end), f"stream {tag} {nz}[{start},{end})")
    nz = 6001
    host = (1.0 + 0.1 * rng.standard_normal((nz, ny, nx))).astype(np.float32)
    host[::5] = -host[::5]
    cube = gpu.Image.upload(host)
    for start, end in [(0, nz), (1, nz)]:
        assert_bit_equal(gpu.call("median", cube, start, end).download(), orc.median(host, start, end), f"stream long[{start},{end})")
    # NaN anywhere in the window -> the reference panics
    host[4000, 50, 60] = np.nan
    e = gpu.get_transform("median").call_results(gpu.Image.upload(host), 0, nz)[0]
    assert isinstance(e, gpu.AflakError) and e.status == gpu.ERR_PANIC


def _median_one_pass_patterns(rng, nz, ny, nx):
    """The patterns above plus the ones that stress a bracket taken from 256 samples: constant spectra (every sample equal:
    the bracket collapses onto the `== lo` class), blank spaxels (all zeros, both signs), two-valued spectra (lo and hi a
    gap apart), a median sitting on +inf, and a spectrum whose upper half is one repeated value (no sample above lo)."""
    out = _median_patterns(rng, nz, ny, nx)
    shape = (nz, ny, nx)
    h = (1.0 + 0.1 * rng.standard_normal(shape)).astype(np.float32)
    h[:, : ny // 3, :] = np.float32(2.5)            # constant spaxels
    h[:, ny // 3 : ny // 2, :] = 0.0                # blank spaxels ...
    h[::2, ny // 3 : ny // 2, ::2] = -0.0           # ... with zeros of both signs
    out.append(("constant and blank spaxels", h))
    h = np.where(rng.random(shape) < 0.5, np.float32(1.0), np.float32(7.0)).astype(np.float32)
    out.append(("two values", h))
    h = (1.0 + 0.1 * rng.standard_normal(shape)).astype(np.float32)
    h[rng.random(shape) < 0.55] = np.inf
    out.append(("median on +inf", h))
    h = rng.uniform(0.0, 1.0, shape).astype(np.float32)
    h[rng.random(shape) < 0.52] = np.float32(1.0)
    out.append(("upper half tied", h))
    return out


def test_median_one_pass_paths(gpu, orc):
    """One-pass streaming select (median_stream2_kernel: 3000..4400 frames on images of >= 9472 spaxels): bracket from 256
    samples, private list segments, wing / open-bin re-collects, refinement of an overflowing list.  Even and odd windows."""
    rng = np.random.default_rng(0xAF1A0043)
    ny, nx = 97, 99
    for nz in (3001, 4400):
        for tag, host in _median_one_pass_patterns(rng, nz, ny, nx):
            cube = gpu.Image.upload(host)
            for start, end in [(0, nz), (1, nz)] if nz < 4400 else [(0, nz)]:
                assert_bit_equal(gpu.call("median", cube, start, end).download(), orc.median(host, start, end), f"one-pass {tag} {nz}[{start},{end})")
            del cube
    host = (1.0 + 0.1 * rng.standard_normal((3200, ny, nx))).astype(np.float32)
    host[1234, 50, 60] = np.nan
    e = gpu.get_transform("median").call_results(gpu.Image.upload(host), 0, 3200)[0]
    assert isinstance(e, gpu.AflakError) and e.status == gpu.ERR_PANIC


def test_median_one_pass_short_windows_subprocess():
    """The same kernel forced onto windows from 801 frames (AFK_MEDIAN_SINGLE_PASS=2; the switch is read once per process)."""
    import os
    import subprocess
    import sys

    here = os.path.dirname(os.path.abspath(__file__))
    env = dict(os.environ, AFK_MEDIAN_SINGLE_PASS="2")
    r = subprocess.run([sys.executable, os.path.join(here, "median_onepass_worker.py")], env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    # a bracket of +-0.5 sigma instead of +-3: most spaxels miss it or end on its last key (re-collect paths, successor above the bracket)
    env = dict(os.environ, AFK_MEDIAN_SINGLE_PASS="2", AFK_MEDIAN_Z="0.5", MEDIAN_WORKER_NZ="1300,2100")
    r = subprocess.run([sys.executable, os.path.join(here, "median_onepass_worker.py")], env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]


@pytest.mark.parametrize("shape", SHAPES)
@pytest.mark.parametrize("is_min", [False, True])
@pytest.mark.parametrize("pathology", [False, True])
def test_argminmax_wavelength_bit_exact(gpu, orc, shape, is_min, pathology):
    nz, ny, nx = shape
    host = orc.gen_cube(0xAF1A0017, nz, ny, nx, pathology=pathology)
    host[nz // 2] = host[nz // 4]  # ties: the first extreme must win
    cube = with_wave_axis(gpu, host)
    for start, end in ranges(nz) + [(3, 4)]:
        arg, val = gpu.call("extract_argmin_max_wavelength", cube, start, end, is_min)
        assert_bit_equal(arg.download(), orc.argmap(host, start, end, 0, is_min, wcs=WCS), f"argmap{shape}[{start},{end})")
        assert_bit_equal(val.download(), orc.minmax_fold(host, start, end, is_min), f"fold{shape}[{start},{end})")
    # result carries the cube's meta unchanged, unit Ang; the fold goes through make_slice
    assert arg.unit == "Ang" and arg.axis(2) == ("WAVE", "Angstrom")
    assert val.unit == "1E-17 erg/s/cm^2/Ang" and val.axis(2) == ("", "")


def test_argmap_axis_selection_and_range(gpu, orc):
    nz, ny, nx = 120, 9, 12
    host = orc.gen_cube(0xAF1A0018, nz, ny, nx)
    # no axis named WAVE: trypix = 0 -> the reference converts the frame index with the WCS of FITS axis 0
    cube = gpu.Image.upload(host)
    cube.set_axis(0, "RA---TAN", "deg")
    cube.set_wcs_axis(0, 10.0, 5.0, -0.25)
    arg, _ = gpu.call("extract_argmin_max_wavelength", cube, 10, 100, False)
    assert_bit_equal(arg.download(), orc.argmap(host, 10, 100, 0, False, wcs=(10.0, 5.0, -0.25)), "argmap trypix=0")
    assert arg.unit == ""
    # meta without WCS keys: pixel index itself
    cube2 = gpu.Image.upload(host)
    cube2.set_axis(2, "WAVE", "Ang")
    arg, _ = gpu.call("extract_argmin_max_wavelength", cube2, 10, 100, True)
    assert_bit_equal(arg.download(), orc.argmap(host, 10, 100, 0, True), "argmap no wcs")
    # peak_based_wavelength_range: indices, never world coordinates, offsets -range / +range (may leave [0, nz))
    cube3 = with_wave_axis(gpu, host)
    for rng in [0, 1, 7, 200]:
        lo, hi = gpu.call("peak_based_wavelength_range", cube3, 0, nz, rng, False)
        assert_bit_equal(lo.download(), orc.argmap(host, 0, nz, -rng, False), f"peak range lo {rng}")
        assert_bit_equal(hi.download(), orc.argmap(host, 0, nz, rng, False), f"peak range hi {rng}")
    # no metadata at all: image.meta().clone().unwrap() panics
    e = gpu.get_transform("extract_argmin_max_wavelength").call_results(gpu.Image.upload(host), 0, 10, False)
    assert all(isinstance(x, gpu.AflakError) and x.status == gpu.ERR_PANIC for x in e)


@pytest.mark.parametrize("shape", SHAPES)
@pytest.mark.parametrize("wcs", [WCS, None])
def test_centroid(gpu, orc, shape, wcs):
    nz, ny, nx = shape
    host = orc.gen_cube(0xAF1A0019, nz, ny, nx)
    cube = with_wave_axis(gpu, host, wcs)
    for start, end in ranges(nz):
        got = gpu.call("extract_centrobaric_wavelength", cube, start, end).download()
        assert_rel(got, orc.centroid_f64(host, start, end, wcs), 1e-5, f"centroid{shape}[{start},{end})")
        gpu.set_sum_mode(gpu.SUM_SEQUENTIAL)
        try:
            got = gpu.call("extract_centrobaric_wavelength", cube, start, end).download()
        finally:
            gpu.set_sum_mode(gpu.SUM_COMPENSATED)
        assert_bit_equal(got, orc.centroid(host, start, end, wcs), f"centroid seq{shape}[{start},{end})")
    res = gpu.call("extract_centrobaric_wavelength", cube, 0, nz)
    assert res.unit == "Ang" and res.has_meta and res.tag is None


def test_centroid_pathology_positions(gpu, orc):
    host = orc.gen_cube(0xAF1A001A, 150, 20, 24, pathology=True)
    cube = with_wave_axis(gpu, host)
    gpu.set_sum_mode(gpu.SUM_SEQUENTIAL)
    try:
        got = gpu.call("extract_centrobaric_wavelength", cube, 0, 150).download()
    finally:
        gpu.set_sum_mode(gpu.SUM_COMPENSATED)
    assert_bit_equal(got, orc.centroid(host, 0, 150, WCS), "centroid pathology")
    got = gpu.call("extract_centrobaric_wavelength", cube, 0, 150).download()
    assert np.array_equal(np.isnan(got), np.isnan(orc.centroid_f64(host, 0, 150, WCS)))


@pytest.mark.parametrize("shape", [(64, 8, 8), (200, 49, 49), (131, 37, 52)])
def test_centroid_with_mask(gpu, orc, shape):
    nz, ny, nx = shape
    host = orc.gen_cube(0xAF1A001B, nz, ny, nx)
    cube = with_wave_axis(gpu, host)
    rng = np.random.default_rng(5)
    lo = rng.uniform(0, nz / 2, (ny, nx)).astype(np.float32)
    hi = (lo + rng.uniform(0, nz / 2, (ny, nx))).astype(np.float32)
    lo[0, 0], hi[0, 0] = 3.0, 3.0          # empty range -> 0/0
    lo[0, 1] = np.nan                       # NaN as usize == 0
    lo[0, 2] = -4.0                         # negative saturates to 0
    hi[-1, -1] = float(nz)                  # end == nz is legal
    mlo, mhi = gpu.Image.upload(lo), gpu.Image.upload(hi)
    gpu.set_sum_mode(gpu.SUM_SEQUENTIAL)
    try:
        got = gpu.call("extract_centrobaric_wavelength_with_mask", cube, mlo, mhi)
    finally:
        gpu.set_sum_mode(gpu.SUM_COMPENSATED)
    assert_bit_equal(got.download(), orc.centroid_mask(host, lo, hi, WCS), f"centroid mask{shape}")
    assert got.unit == "" and not got.has_meta
    want64 = orc.centroid_mask(host, lo, hi, WCS).astype(np.float64)
    got = gpu.call("extract_centrobaric_wavelength_with_mask", cube, mlo, mhi).download()
    fin = np.isfinite(want64)
    assert np.array_equal(np.isnan(got), np.isnan(want64))
    assert np.all(np.abs(got[fin] - want64[fin]) <= 2e-4 * np.abs(want64[fin]))  # want64 is the f32 sequential result
    # end beyond the cube or start > end: slice_axis panics
    hi_bad = hi.copy()
    hi_bad[1, 1] = nz + 1
    e = gpu.get_transform("extract_centrobaric_wavelength_with_mask").call_results(cube, mlo, gpu.Image.upload(hi_bad))[0]
    assert isinstance(e, gpu.AflakError) and e.status == gpu.ERR_PANIC
    e = gpu.get_transform("extract_centrobaric_wavelength_with_mask").call_results(cube, mhi, mlo)[0]
    assert isinstance(e, gpu.AflakError) and e.status == gpu.ERR_PANIC
    # mask of another shape
    e = gpu.get_transform("extract_centrobaric_wavelength_with_mask").call_results(cube, gpu.Image.upload(lo[:, :-1]), mhi)[0]
    assert isinstance(e, gpu.AflakError) and e.status == gpu.ERR_PANIC


def test_pix2world_abi(gpu, orc):
    img = gpu.Image.upload(np.zeros((4, 3, 2), np.float32))
    assert img.pix2world(2, 5.0) is None  # no metadata: Option::None
    img.set_axis(2, "WAVE", "Ang")
    assert img.pix2world(2, 5.0) == 5.0
    img.set_wcs_axis(2, *WCS)
    for p in [0.0, 1.0, 5.0, 4562.0, -3.0]:
        assert img.pix2world(2, p) == orc.pix2world(WCS, p)
    assert img.pix2world(0, 7.0) == 7.0  # other axes stay identity
