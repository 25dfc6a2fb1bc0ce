"""CPU: pin the oracle (oracle/, the parity checker) against the committed fixtures of tests/golden/.

  lut_golden.json   the reference's own value-asserting unit tests (aflak_plot/src/imshow/lut.rs:292-343)
  kat.json          hand-derived known answers for operators the reference never asserts values for
  test_fits_*       a crop of the reference's data fixture with numpy-f64 check sums
"""
import json
import os

import numpy as np
import pytest

from util import assert_bit_equal

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    with open(os.path.join(GOLD, name)) as f:
        return json.load(f)


def grad(g):
    return [(s[0], tuple(s[1])) for s in g]


def test_lut_color_at_golden(orc):
    for case in load("lut_golden.json")["color_at"]:
        table = orc.lut_build(grad(case["gradient"]), tuple(case.get("table_lims", case["lims"])))
        for point, rgb in case["cases"]:
            assert orc.color_at(table, point, tuple(case["lims"])) == rgb, (point, rgb)


def test_lut_bounds_golden(orc):
    b = load("lut_golden.json")["bounds"]
    got = orc.lut_bounds(grad(b["gradient"]), tuple(b["lims"]))
    want = [((p[0][0], tuple(p[0][1])), (p[1][0], tuple(p[1][1]))) for p in b["pairs"]]
    assert got == want


def test_lut_builtin_tables_are_monotone_in_grey(orc):
    t = orc.lut_build("Grey")
    assert t.shape == (65536, 3)
    assert np.all(np.diff(t[:, 0].astype(np.int32)) >= 0)
    assert tuple(t[0]) == (0, 0, 0) and tuple(t[-1]) == (255, 255, 255)
    assert tuple(t[32767]) == (127, 127, 127)  # truncating `as u8`: 255 * 0.49999 = 127.49 -> 127


def test_integral_ramp_kat(orc):
    k = load("kat.json")
    for c in k["integral_ramp"]:
        nz = c["nz"]
        cube = np.broadcast_to(np.arange(nz, dtype=np.float32)[:, None, None], (nz, 3, 5)).copy()
        out = orc.integral(cube, c["a"], c["b"])
        assert np.all(out == np.float32(c["value"]))
        avg = orc.average(cube, c["a"], c["b"])
        assert np.all(avg == np.float32(np.float32(c["value"]) / np.float32(c["b"] - c["a"])))
    for c in k["middle_width"]:
        assert orc.middle_width(c["a"], c["b"]) == (c["middle"], c["width"])


def test_elementwise_kat(orc):
    k = load("kat.json")
    for c in k["equivalent_width"]:
        out = orc.equivalent_width(np.array([c["off"]], np.float32), np.array([c["on"]], np.float32), c["fl"], c["max"], c["emission"])
        assert out[0] == np.float32(c["value"]), c
    r = k["ratio_from_bands"]
    r2, r1 = orc.ratio_from_bands(r["z"], r["z1"], r["z2"])
    assert (np.float32(r2), np.float32(r1)) == (np.float32(r["off_ratio_2"]), np.float32(r["off_ratio_1"]))
    for c in k["linear_composition"]:
        out = orc.linear_composition(np.array([c["u"]], np.float32), np.array([c["v"]], np.float32), c["a"], c["b"])
        assert out[0] == np.float32(c["out"])
    for c in k["logscale"]:
        out = orc.convert_to_logscale(np.array([c["v"]], np.float32), c["a"], c["vmin"], c["vmax"])
        assert out[0] == np.float32(c["out"])
    table = orc.tone_curve_default()
    assert table.size == 256 and table[255] == 255.0
    for c in k["tone_curve_default"]:
        out = orc.apply_tone_curve(np.array([c["v"]], np.float32), table)
        assert out[0] == np.float32(c["out"]), c
    mn, mx = orc.image_min_max(np.zeros((0,), np.float32))
    assert (float(mn), float(mx)) == (k["image_min_max_empty"]["min"], k["image_min_max_empty"]["max"])


def test_down_sampling_kat(orc):
    d = load("kat.json")["down_sampling"]
    img = np.arange(np.prod(d["shape"]), dtype=np.float32).reshape(d["shape"])
    out = orc.down_sampling(img, d["n"])
    assert list(out.shape) == d["out_shape"]
    want = np.array([img[y, x] for y, x in d["picks"]], np.float32).reshape(d["out_shape"])
    assert_bit_equal(out, want)
    cube = np.arange(2 * 4 * 5, dtype=np.float32).reshape(2, 4, 5)
    out3 = orc.down_sampling(cube, 3)
    assert out3.shape == (2, 2, 2)
    assert_bit_equal(out3, cube[:, ::3, ::3])


def test_error_strings(orc):
    e = load("kat.json")["errors"]
    cube = np.zeros((10, 2, 2), np.float32)
    cases = [(lambda: orc.integral(cube, -1, 5), e["start_negative"]), (lambda: orc.integral(cube, 2, -3), e["end_negative"]),
             (lambda: orc.integral(cube, 5, 3), e["start_ge_end"]), (lambda: orc.average(cube, 0, 11), e["end_too_large_nz10"]),
             (lambda: orc.slice_one_frame(cube, 10), e["frame_too_large_nz10"]),
             (lambda: orc.down_sampling(cube[0], 0), e["down_sampling_n0"]),
             (lambda: orc.ratio_from_bands(3.0, 5.0, 9.0), e["ratio_from_bands_order"])]
    for fn, msg in cases:
        with pytest.raises(orc.OracleError) as ei:
            fn()
        assert ei.value.msg == msg


def test_real_cube_crop(orc):
    """The reference's data fixture (crop): the f32 sequential oracle vs numpy f64 sums, and bit-exact slicing."""
    crop = np.load(os.path.join(GOLD, "test_fits_crop.npy"))
    facts = load("test_fits_facts.json")
    assert crop.shape == (200, 12, 12) and crop.dtype == np.float32
    want = np.array(facts["crop_sum_f64_40_160"])
    got = orc.integral(crop, 40, 160).astype(np.float64)
    np.testing.assert_allclose(got, want, rtol=2e-5, atol=1e-4)  # the reference's own f32 error: ~n * 2^-24 relative
    s64, a64 = orc.sum_f64(crop, 40, 160)
    np.testing.assert_allclose(s64, want, rtol=1e-12)
    assert_bit_equal(orc.slice_one_frame(crop, 100), np.array(facts["crop_frame_100"], np.float32))


def test_sequential_sum_order(orc):
    """ndarray 0.12.1 sum_axis: strictly left-to-right f32 per pixel -- differs from pairwise/f64 on purpose."""
    col = np.array([1e8, 1.0, -1e8, 1.0], np.float32).reshape(4, 1, 1)
    assert orc.integral(col, 0, 4)[0, 0] == np.float32(1.0)  # (1e8 + 1) rounds to 1e8; then -1e8 -> 0; + 1 -> 1
    s64, _ = orc.sum_f64(col, 0, 4)
    assert s64[0, 0] == 2.0


def test_average_count_is_f32_division(orc):
    cube = np.full((3, 1, 1), 0.1, np.float32)
    s = np.float32(np.float32(np.float32(0.1) + np.float32(0.1)) + np.float32(0.1))
    assert orc.average(cube, 0, 3)[0, 0] == np.float32(s / np.float32(3.0))


def test_generator_is_deterministic_and_shardable(orc):
    full = orc.gen_cube(7, 9, 10, 6)
    band = orc.gen_cube(7, 9, 10, 6, y0=4, rows=3)
    assert_bit_equal(band, full[:, 4:7, :])
    assert np.all(full > 0)
    p = orc.gen_cube(7, 64, 32, 32, pathology=True)
    assert np.isnan(p).any() and np.isinf(p).any()


def test_get_vmin_vmax_skip_nan(orc):
    a = np.array([np.nan, 3.0, -2.0, np.nan, 7.0], np.float32)
    assert orc.get_vmin(a) == np.float32(-2.0) and orc.get_vmax(a) == np.float32(7.0)
    assert np.isnan(orc.get_vmin(np.array([np.nan, np.nan], np.float32)))
    mn, mx = orc.image_min_max(np.array([np.nan, np.inf, -1.0], np.float32))
    assert mn == np.float32(-1.0) and np.isinf(mx)


def test_make_raw_image_flip_and_nan(orc):
    img = np.array([[0.0, 1.0], [np.nan, 0.5]], np.float32)
    table = orc.lut_build("Grey")
    out = orc.make_raw_image(img, 0.0, 1.0, table, channels=3)
    # output row 0 is input row 1 (vertical flip, image.rs:123-129); NaN -> index 0 -> lut[0]
    assert out[0, 0].tolist() == [0, 0, 0] and out[0, 1].tolist() == [127, 127, 127]
    assert out[1, 0].tolist() == [0, 0, 0] and out[1, 1].tolist() == [255, 255, 255]
    rgba = orc.make_raw_image(img, 0.0, 1.0, table, channels=4)
    assert np.all(rgba[..., 3] == 255) and np.array_equal(rgba[..., :3], out)


# ---- other axis-0 reductions (lib.rs:1332-1582): hand-derivable known answers ------------------------------------

def test_variance_stddev_kat(orc):
    # spectrum 1,2,3,4 per pixel: mean 2.5, sum of squares 5, ddof=1 -> 5/3; every Welford step is exact in f32 but the last division
    cube = np.arange(1, 5, dtype=np.float32).reshape(4, 1, 1) * np.ones((1, 2, 3), np.float32)
    v = orc.variance(cube, 0, 4)
    assert v.shape == (2, 3) and np.all(v == np.float32(5.0) / np.float32(3.0))
    assert np.all(orc.stddev(cube, 0, 4) == np.sqrt(np.float32(5.0) / np.float32(3.0), dtype=np.float32))
    # constant spectrum -> exactly 0; two frames a,b -> (a-b)^2/2
    assert np.all(orc.variance(np.full((9, 2, 2), 3.25, np.float32), 0, 9) == 0.0)
    two = np.array([1.0, 4.0], np.float32).reshape(2, 1, 1)
    assert orc.variance(two, 0, 2)[0, 0] == 4.5
    # ndarray panics when ddof >= number of frames
    with pytest.raises(orc.OracleError) as e:
        orc.variance(cube, 1, 2)
    assert e.value.code == orc.ERR_PANIC
    # agrees with numpy's two-pass f64 variance on real data
    crop = np.load(os.path.join(GOLD, "test_fits_crop.npy"))
    v64 = crop[40:160].astype(np.float64).var(axis=0, ddof=1)
    assert np.allclose(orc.variance_f64(crop, 40, 160), v64, rtol=1e-12)
    assert np.allclose(orc.variance(crop, 40, 160), v64, rtol=2e-4)


def test_median_kat(orc):
    cube = np.array([5, 1, 4, 2, 3], np.float32).reshape(5, 1, 1)
    assert orc.median(cube, 0, 5)[0, 0] == 3.0      # odd: the middle element
    assert orc.median(cube, 0, 4)[0, 0] == 3.0      # even: (2 + 4) / 2
    assert orc.median(cube, 1, 3)[0, 0] == 2.5
    assert orc.median(cube, 2, 3)[0, 0] == 4.0
    bad = cube.copy()
    bad[1] = np.nan
    with pytest.raises(orc.OracleError) as e:       # partial_cmp().unwrap() on NaN
        orc.median(bad, 0, 5)
    assert e.value.code == orc.ERR_PANIC
    assert np.isnan(orc.median(bad, 1, 2)[0, 0])    # a single frame is never compared


def test_argmap_and_fold_kat(orc):
    spec = np.array([1, 7, 3, 7, np.nan, 2], np.float32).reshape(6, 1, 1)
    # first maximum wins (strict >), NaN never wins
    assert orc.argmap(spec, 0, 6, 0, False)[0, 0] == 1.0
    assert orc.argmap(spec, 2, 6, 0, False)[0, 0] == 3.0
    assert orc.argmap(spec, 0, 6, 0, True)[0, 0] == 0.0
    assert orc.argmap(spec, 0, 6, -2, False)[0, 0] == -1.0   # peak_based_wavelength_range: index - range
    assert orc.argmap(spec, 4, 5, 0, False)[0, 0] == 0.0     # nothing wins -> 0.0
    # linear WCS: CRVAL 4000 at CRPIX 1, CDELT 0.5 -> pixel 1 -> 4000.5
    assert orc.argmap(spec, 0, 6, 0, False, wcs=(4000.0, 1.0, 0.5))[0, 0] == 4000.5
    # the fold restarts after a NaN: max over frames after the last NaN
    assert orc.minmax_fold(spec, 0, 6, False)[0, 0] == 2.0
    assert orc.minmax_fold(spec, 0, 4, False)[0, 0] == 7.0
    assert np.isnan(orc.minmax_fold(spec, 0, 5, False)[0, 0])
    assert orc.minmax_fold(spec, 0, 4, True)[0, 0] == 1.0


def test_centroid_kat(orc):
    # flux 1,2,1 at pixels 10,11,12 -> centroid 11 exactly
    cube = np.zeros((20, 1, 2), np.float32)
    cube[10:13, 0, 0] = [1, 2, 1]
    cube[10:13, 0, 1] = [1, 0, 1]
    c = orc.centroid(cube, 8, 15)
    assert c[0, 0] == 11.0 and c[0, 1] == 11.0
    assert orc.centroid(cube, 8, 15, wcs=(100.0, 1.0, 2.0))[0, 0] == 122.0
    assert np.isnan(orc.centroid(cube, 0, 5)[0, 0])          # 0 / 0
    lo = np.array([[10.9, 0.0]], np.float32)                  # `as usize` truncates
    hi = np.array([[13.0, 20.0]], np.float32)
    m = orc.centroid_mask(cube, lo, hi)
    assert m[0, 0] == 11.0 and m[0, 1] == 11.0
    with pytest.raises(orc.OracleError):
        orc.centroid_mask(cube, lo, np.array([[21.0, 20.0]], np.float32))


# ---- display statistics (aflak_plot hist.rs, lims.rs:37-106, zscale.rs): known answers ---------------------------------

def test_histogram_kat(orc):
    img = np.array([[0.0, 0.5, 9.99, 10.0], [np.nan, -3.0, 25.0, 5.0]], np.float32)
    counts, starts, ends = orc.histogram(img, 0.0, 10.0)
    assert counts.sum() == 8                      # every pixel is counted
    assert counts[0] == 3                         # 0.0, NaN (as usize -> 0) and -3.0 (saturates to 0)
    assert counts[5] == 1 and counts[50] == 1     # 0.5 -> 5, 5.0 -> 50
    assert counts[99] == 3                        # 9.99 -> 99, 10.0 -> 100 -> clamped, 25.0 clamped
    assert starts[0] == 0.0 and ends[99] == np.float32(0.0) + np.float32(100.0) * (np.float32(10.0) / np.float32(100.0))
    rgb = np.stack([np.full((2, 2), 0.0), np.full((2, 2), 65535.0), np.full((2, 2), 32768.0)]).astype(np.float32)
    c3, _, _ = orc.histogram(rgb, 0.0, 65535.0, color=True)
    assert c3.shape == (256, 3) and c3[0, 0] == 4 and c3[255, 1] == 4 and c3[128, 2] == 4 and c3.sum() == 12


def test_vmed_vmad_kat(orc):
    # values 0..4 -> normalised 0, .25, .5, .75, 1 -> median .5, deviations .5 .25 0 .25 .5 -> MAD .25
    img = np.array([[3, 0, 4], [1, 2, np.nan]], np.float32)
    vmed, vmad = orc.vmed_vmad(img)
    assert vmed == 0.5 and vmad == 0.25
    # even count: mean of the two middle values
    img = np.array([0, 1, 2, 4], np.float32)
    vmed, vmad = orc.vmed_vmad(img)
    assert vmed == np.float32(0.375) and vmad == np.float32((np.float32(0.125) + np.float32(0.375)) / 2)
    with pytest.raises(orc.OracleError):          # constant image: (v - vmin) / 0 is NaN everywhere
        orc.vmed_vmad(np.full((3, 3), 2.0, np.float32))


def test_zscale_kat(orc):
    # a linear ramp is its own fit: nothing is rejected.  30x30 -> stride 1, all 900 samples, slope 1 per sample:
    # z1 = max(zmin, median - (center_pixel - 1) * slope / contrast), z2 = min(zmax, median + (npix - center_pixel) * slope / contrast)
    img = np.arange(30 * 30, dtype=np.float32).reshape(30, 30)
    assert orc.zscale(img, 1000, 0.25) == (0.0, 899.0)            # slope x4: both limits clamp to the sample range
    z1, z2 = orc.zscale(img, 1000, 1.0)
    assert abs(float(z1) - (449.5 - 448.0)) < 1e-2 and z2 == 899.0
    # 32x32 with nsamples 1000: the sample is the first 1000 pixels of the *flipped* image (rows 31..1 and 8 pixels of row 0)
    img = np.arange(32 * 32, dtype=np.float32).reshape(32, 32)
    assert orc.zscale(img, 1000, 0.25) == (0.0, 1023.0)
    img[0, :8] = -50.0   # inside the sample
    img[0, 8:] = -99.0   # beyond the truncation
    assert orc.zscale(img, 1000, 0.25)[0] == -50.0
    assert orc.zscale(np.full((4, 4), np.nan, np.float32)) == (None, None)
    assert orc.zscale(img, 0, 0.25) == (None, None)
    z1, z2 = orc.zscale(np.full((40, 40), 3.0, np.float32))   # constant image: slope 0 -> both limits are the value
    assert z1 == 3.0 and z2 == 3.0


# ---- spatial reductions / gathers and remaining elementwise operators: known answers ------------------------------------

def test_extract_wave_and_extrude_kat(orc):
    cube = np.arange(3 * 4 * 4, dtype=np.float32).reshape(3, 4, 4)
    # ROI::All on a square frame: plain sum of the frame
    assert orc.extract_wave(cube).tolist() == [120.0, 376.0, 632.0]
    # PixelList (i, j) = (column, row); pixels outside the frame are skipped
    assert orc.extract_wave(cube, [(1, 0), (3, 2), (9, 9)]).tolist() == [1.0 + 11.0, 17.0 + 27.0, 33.0 + 43.0]
    assert orc.extract_wave(cube, []).tolist() == [0.0, 0.0, 0.0]
    with pytest.raises(orc.OracleError) as e:   # the swapped loop bounds of ROI::All index out of range on non-square frames
        orc.extract_wave(np.zeros((2, 3, 4), np.float32))
    assert e.value.code == orc.ERR_PANIC
    # extrude: frames in reverse order, one column per pixel
    ex = orc.extrude(cube, [(1, 0), (3, 2)])
    assert ex.tolist() == [[33.0, 43.0], [17.0, 27.0], [1.0, 11.0]]
    for bad in (None, [(1, 0), (9, 9)]):       # ROI::All has datalen 0; a skipped pixel leaves the vector short
        with pytest.raises(orc.OracleError):
            orc.extrude(cube, bad)


def test_plane_and_slice_kat(orc):
    cube = np.arange(5 * 6 * 7, dtype=np.float32).reshape(5, 6, 7)
    # defaults of make_plane3d: dir1 = z, dir2 = y; p0 = (2, 0, 0) -> the frame cube[2]
    xyz = orc.make_plane3d((2, 0, 0), (0, 0, 1), (0, 1, 0), 7, 6)
    assert xyz.shape == (6, 7, 3) and xyz[3, 4].tolist() == [2.0, 3.0, 4.0]
    assert np.array_equal(orc.slice_3d_to_2d(cube, xyz), cube[2])
    # nearest voxel by truncation; negative and NaN coordinates become index 0
    pts = np.array([[[1.9, 2.2, 3.99], [-0.5, np.nan, 0.0]]], np.float32)
    assert orc.slice_3d_to_2d(cube, pts).tolist() == [[cube[1, 2, 3], cube[0, 0, 0]]]
    with pytest.raises(orc.OracleError) as e:
        orc.slice_3d_to_2d(cube, np.array([[[0, 0, 0], [5, 0.5, 2]]], np.float32))
    assert e.value.msg == "Input maps to out of bound pixel!: [5, 0.5, 2]"


def test_elementwise_next_kat(orc):
    # velocity: (w - w0) * 3e5 / w0
    assert orc.velocity_field(np.array([6562.8, 6569.3628], np.float32), 6562.8)[0] == 0.0
    v = orc.velocity_field(np.array([2.0], np.float32), 1.0)
    assert v[0] == 300000.0
    # histogram transformation with the default parameters (0, 0.5, 1) is the identity up to rounding: c = x
    img = np.array([0.0, 1.0, 2.0, 4.0], np.float32)
    assert orc.histogram_transformation(img, (0.0, 0.5, 1.0)).tolist() == [0.0, 1.0, 2.0, 4.0]
    # midtone 0.25 lifts the middle: x = .5 -> (m-1)x / ((2m-1)x - m) = -.375 / -.5 = .75
    assert orc.histogram_transformation(img, (0.0, 0.25, 1.0))[2] == 3.0
    # clipping: x > 1 -> max, x < 0 -> min
    out = orc.histogram_transformation(img, (0.25, 0.5, 0.5))
    assert out[0] == 0.0 and out[3] == 4.0 and out[2] == 4.0
    # hsv of pure colours (scaled to 65535)
    rgb = np.array([[[1.0, 0.0, 0.0, 0.5]], [[0.0, 1.0, 0.0, 0.5]], [[0.0, 0.0, 1.0, 0.5]]], np.float32)
    hsv = orc.color_image_to_hsv(rgb)
    assert hsv[0, 0, 0] == 0.0 and hsv[0, 0, 1] == np.float32(120.0) / np.float32(360.0) * np.float32(65535.0)
    assert hsv[1, 0, 0] == 65535.0 and hsv[2, 0, 2] == 1.0
    assert np.isnan(hsv[0, 0, 3]) and hsv[1, 0, 3] == 0.0   # grey: 0/0 hue, zero saturation
    # arcsinh stretch with a tiny beta is the identity to first order: asinh(l*b) / (l * asinh(b)) -> 1
    out = orc.arcsinh_stretch(np.full((3, 1, 2), 100.0, np.float32), 1e-9)
    assert np.allclose(out, 100.0, rtol=1e-6)
