"""GPU parity on the reference's own data fixture, whole: tests/golden/test.fits is
aflak_primitives/test/test.fits (200 x 49 x 49, BITPIX -32), opened through the registry exactly like the reference's
test does (aflak_primitives/src/lib.rs:1969-1991: open_fits -> fits_to_image -> operators), then every operator of the hot
path is run on it and compared with the C oracle AND with the independent numpy twin (tests/twin.py).
Sequential sum mode: bit for bit.  Default (compensated) mode: tier C against the f64 sum.
"""
import os

import numpy as np
import pytest

import twin
from util import assert_bit_equal, assert_reduction, assert_ulp

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
# FITS axis 0 of the file (CRVAL1, CRPIX1, CDELT1): no axis is named WAVE, so run_create_argmap / run_centroid convert the
# frame index with the WCS of axis `trypix = 0` (lib.rs:1408-1418, 1477-1487)
WCS0 = (-6.000000E-02, 2.499996E+01, -2.083333E-03)
WINDOWS = [(0, 200), (40, 160), (17, 18), (190, 200)]


@pytest.fixture(scope="module")
def fixture_cube(gpu):
    fits = gpu.call("open_fits", gpu.Paths([os.path.join(GOLD, "test.fits")]), 0)
    cube = gpu.call("fits_to_image", fits, 0, "")
    raw = open(os.path.join(GOLD, "test.fits"), "rb").read()
    host = np.frombuffer(raw[2880:2880 + 200 * 49 * 49 * 4], dtype=">f4").astype(np.float32).reshape(200, 49, 49)
    return cube, host


def test_decode_and_metadata(gpu, fixture_cube):
    cube, host = fixture_cube
    assert cube.shape == (200, 49, 49)
    assert_bit_equal(cube.download(), host, "fits_to_image(test.fits)")
    assert cube.unit == "K" and cube.axis(2)[0] == "VELO-LSR" and cube.axis(0)[0] == "GLON-GLS"
    w = cube.pix2world(0, 0.0)
    assert w is not None and np.float32(w) == np.float32(WCS0[0] + WCS0[2] * (0.0 + 1.0 - WCS0[1]))


def test_reductions_sequential_bit_exact(gpu, orc, fixture_cube):
    cube, host = fixture_cube
    gpu.set_sum_mode(gpu.SUM_SEQUENTIAL)
    try:
        for s, e in WINDOWS:
            img, mid, width = gpu.call("integral", cube, s, e)
            assert_bit_equal(img.download(), orc.integral(host, s, e), f"integral[{s},{e}) vs oracle")
            assert_bit_equal(img.download(), twin.sum_axis(host, s, e), f"integral[{s},{e}) vs twin")
            assert (float(mid), float(width)) == orc.middle_width(s, e)
            avg = gpu.call("average", cube, s, e)[0].download()
            assert_bit_equal(avg, orc.average(host, s, e), f"average[{s},{e}) vs oracle")
            assert_bit_equal(avg, twin.mean_axis(host, s, e), f"average[{s},{e}) vs twin")
            if e - s >= 2:
                assert_bit_equal(gpu.call("variance", cube, s, e).download(), twin.var_axis(host, s, e), f"variance[{s},{e}) vs twin")
                assert_bit_equal(gpu.call("stddev", cube, s, e).download(), orc.stddev(host, s, e), f"stddev[{s},{e}) vs oracle")
            c = gpu.call("extract_centrobaric_wavelength", cube, s, e).download()
            assert_bit_equal(c, orc.centroid(host, s, e, wcs=WCS0), f"centroid[{s},{e}) vs oracle")
            world = lambda p: np.float32(WCS0[0] + WCS0[2] * (float(p) + 1.0 - WCS0[1]))
            assert_bit_equal(c, twin.centroid(host, s, e, world=world), f"centroid[{s},{e}) vs twin")
    finally:
        gpu.set_sum_mode(gpu.SUM_COMPENSATED)


def test_reductions_default_mode_tier_c(gpu, orc, fixture_cube):
    cube, host = fixture_cube
    for s, e in WINDOWS:
        s64, a64 = orc.sum_f64(host, s, e)
        assert_reduction(gpu.call("integral", cube, s, e)[0].download(), s64, a64, f"integral[{s},{e})")
        assert_reduction(gpu.call("average", cube, s, e)[0].download(), s64, a64, f"average[{s},{e})", divisor=float(e - s))
        if e - s >= 2:
            v64 = orc.variance_f64(host, s, e)
            got = gpu.call("variance", cube, s, e).download().astype(np.float64)
            assert np.all(np.abs(got - v64) <= 1e-5 * np.abs(v64) + 1e-12), f"variance[{s},{e})"


def test_order_statistics_bit_exact(gpu, orc, fixture_cube):
    cube, host = fixture_cube
    for s, e in WINDOWS:
        med = gpu.call("median", cube, s, e).download()
        assert_bit_equal(med, orc.median(host, s, e), f"median[{s},{e}) vs oracle")
        assert_bit_equal(med, twin.median(host, s, e), f"median[{s},{e}) vs twin")
        for is_min in (False, True):
            arg, val = gpu.call("extract_argmin_max_wavelength", cube, s, e, is_min)
            assert_bit_equal(arg.download(), orc.argmap(host, s, e, 0, is_min, wcs=WCS0), f"argmap[{s},{e}) min={is_min}")
            assert_bit_equal(val.download(), twin.fold_minmax(host, s, e, is_min), f"fold[{s},{e}) min={is_min} vs twin")
            lo, hi = gpu.call("peak_based_wavelength_range", cube, s, e, 5, is_min)
            assert_bit_equal(lo.download(), twin.argmap(host, s, e, -5, is_min), "peak range lo vs twin")
            assert_bit_equal(hi.download(), twin.argmap(host, s, e, 5, is_min), "peak range hi vs twin")


def test_elementwise_and_display_bit_exact(gpu, orc, fixture_cube):
    cube, host = fixture_cube
    gpu.set_sum_mode(gpu.SUM_SEQUENTIAL)
    try:
        on, off1, off2 = (90, 110), (40, 60), (140, 160)
        a_on, a1, a2 = (gpu.call("average", cube, *w)[0] for w in (on, off1, off2))
        h_on, h1, h2 = (orc.average(host, *w) for w in (on, off1, off2))
    finally:
        gpu.set_sum_mode(gpu.SUM_COMPENSATED)
    for f in (0, 57, 199):
        assert_bit_equal(gpu.call("slice_one_frame", cube, f).download(), host[f], f"slice_one_frame {f}")
    r2, r1 = gpu.call("ratio_from_bands", 100.0, 50.0, 150.0)
    assert (np.float32(r2), np.float32(r1)) == tuple(np.float32(v) for v in orc.ratio_from_bands(100.0, 50.0, 150.0))
    off = gpu.call("linear_composition", a1, a2, float(r2), float(r1))
    h_off = orc.linear_composition(h1, h2, float(r2), float(r1))
    assert_bit_equal(off.download(), h_off, "linear_composition")
    assert_bit_equal(gpu.call("ratio", a_on, off).download(), orc.ratio(h_on, h_off), "ratio")
    ew = gpu.call("create_equivalent_width", off, a_on, 40.0, 1e10, True)
    h_ew = orc.equivalent_width(h_off, h_on, 40.0, 1e10, True)
    assert_bit_equal(ew.download(), h_ew, "create_equivalent_width")
    fused = gpu.fused_equivalent_width  # the template's fused kernel on the same windows: sequential mode = the same bits
    gpu.set_sum_mode(gpu.SUM_SEQUENTIAL)
    try:
        assert_bit_equal(fused(cube, on, off1, off2, 40.0, 1e10, True).download(), h_ew, "fused EW template")
    finally:
        gpu.set_sum_mode(gpu.SUM_COMPENSATED)
    mn, mx = gpu.call("image_min_max", ew)
    assert (np.float32(mn), np.float32(mx)) == tuple(np.float32(v) for v in orc.image_min_max(h_ew))
    assert_ulp(gpu.call("convert_to_logscale", ew, 1000.0, float(mn), float(mx)).download(), orc.convert_to_logscale(h_ew, 1000.0, mn, mx), 2, "logscale")
    for n in (2, 3, 7):
        assert_bit_equal(gpu.call("down_sampling", a_on, n).download(), orc.down_sampling(h_on, n), f"down_sampling {n}")
    assert_bit_equal(gpu.call("down_sampling", cube, 4).download(), orc.down_sampling(host, 4), "down_sampling 3-D")
    # display: limits, colour bytes, tone curve (aflak_plot/src/lims.rs, imshow/image.rs, lib.rs:1858-1871)
    vmin, vmax = a_on.vmin_vmax()
    assert (np.float32(vmin), np.float32(vmax)) == (np.float32(orc.get_vmin(h_on)), np.float32(orc.get_vmax(h_on)))
    lut = gpu.Lut("Flame")
    for ch in (3, 4):
        got = gpu.colormap(a_on, lut, channels=ch)
        want = orc.make_raw_image(h_on, orc.get_vmin(h_on), orc.get_vmax(h_on), orc.lut_build("Flame"), channels=ch)
        assert np.array_equal(got, want), f"colormap {ch} channels"
    tone = orc.create_curve([(0.0, 0.0), (0.3, 0.6), (1.0, 1.0)])
    nm = gpu.call("normalize_image", a_on, float(vmin), float((vmax - vmin) / 255.0 + vmin))
    h_nm = ((h_on - np.float32(vmin)) / (np.float32((vmax - vmin) / 255.0 + vmin) - np.float32(vmin))).astype(np.float32)
    assert_bit_equal(nm.download(), h_nm, "normalize_image")
    clipped = gpu.call("clip_image", nm, 255.0, True, 0.0, True)
    h_cl = np.where((h_nm >= np.float32(255.0)) | (h_nm <= np.float32(0.0)), np.float32(np.nan), h_nm)  # run_clip writes NaN (lib.rs:1267-1271)
    assert_bit_equal(clipped.download(), h_cl, "clip_image")
    assert_bit_equal(gpu.call("apply_tone_curve", clipped, gpu.ToneCurve(tone)).download(), orc.apply_tone_curve(h_cl, tone), "apply_tone_curve")
    for got, want in zip(a_on.histogram(float(vmin), float(vmax)), orc.histogram(h_on, vmin, vmax)):
        assert np.array_equal(got, want), "histogram"
    assert a_on.vmed_vmad() == orc.vmed_vmad(h_on)
    # spatial gathers over the whole cube
    assert_ulp(gpu.call("extract_wave", cube, gpu.Roi()).download(), orc.extract_wave(host, None, f64=True).astype(np.float32), 4, "extract_wave all")
    px = [(x, y) for y in range(20, 30) for x in range(15, 35)]
    assert_bit_equal(gpu.call("extrude", cube, gpu.Roi(px)).download(), orc.extrude(host, px), "extrude")
