"""GPU parity: spatial reductions / gathers and the remaining elementwise operators.

Reference: aflak_primitives/src/lib.rs:198-225, 360-383, 509-631 (registry), 928-1169, 1233-1256, 1795-1892; roi.rs:12-37.
Tier A (bit-exact): extrude, slice_3d_to_2d, make_plane3d, velocity field, histogram transformation, hsv, colour stacking, and
extract_wave in SUM_SEQUENTIAL mode.  Tier C: extract_wave in the default mode, <= 1e-5 relative to the f64 sum (or the
4 * 2^-24 * sum|x| cancellation floor).  Tier B: arcsinh stretch, <= 1 ulp (libm asinh in f64, rounded to f32).
"""
import numpy as np
import pytest

from util import assert_bit_equal, assert_ulp

pytestmark = pytest.mark.gpu


def cube_with_meta(gpu, host):
    cube = gpu.Image.upload(host)
    cube.unit = "Jy/beam"
    cube.set_axis(0, "RA---SIN", "deg")
    cube.set_axis(1, "DEC--SIN", "deg")
    cube.set_axis(2, "FREQ", "Hz")
    cube.set_wcs_axis(2, 1.0e11, 1.0, -5.0e5)
    cube.tag = "some tag"
    return cube


def rois(rng, ny, nx):
    yield "one pixel", [(nx // 2, ny // 2)]
    yield "box", [(x, y) for y in range(ny // 4, ny // 2) for x in range(nx // 3, nx // 3 + 7)]
    pts = [(int(rng.integers(0, nx)), int(rng.integers(0, ny))) for _ in range(2000)]
    yield "random with repeats", pts
    yield "partly outside", [(nx - 1, ny - 1), (nx, 0), (0, ny), (10 ** 12, 3), (1, 1)]
    yield "empty", []


# (37,64,64), (70,46,46), (33,36,36): whole-frame sequential sums take the staged kernel (ragged last warp, 68- and 16-value tails)
@pytest.mark.parametrize("shape", [(37, 64, 64), (200, 49, 49), (9, 30, 52), (600, 8, 8), (70, 46, 46), (33, 36, 36)])
def test_extract_wave(gpu, orc, shape):
    nz, ny, nx = shape
    host = orc.gen_cube(0xAF1A0031, nz, ny, nx)
    host[::5] *= -1.0
    cube = cube_with_meta(gpu, host)
    rng = np.random.default_rng(31)
    cases = list(rois(rng, ny, nx))
    if ny == nx:
        cases.append(("all", None))
    for name, px in cases:
        roi = gpu.Roi(px)
        got = gpu.call("extract_wave", cube, roi)
        assert got.shape == (nz,)
        w64 = orc.extract_wave(host, px, f64=True)
        if px is None:
            a64 = np.abs(host.astype(np.float64)).sum(axis=(1, 2))
        else:
            inside = [(x, y) for x, y in px if x < nx and y < ny]
            a64 = np.array([sum(abs(float(host[k, y, x])) for x, y in inside) for k in range(nz)]) if len(inside) < 100 else \
                np.abs(host[:, [p[1] for p in inside], [p[0] for p in inside]].astype(np.float64)).sum(axis=1)
        err = np.abs(got.download().astype(np.float64) - w64)
        assert np.all((err <= 1e-5 * np.abs(w64)) | (err <= 4 * 2.0 ** -24 * a64)), f"extract_wave {name} {shape}"
        gpu.set_sum_mode(gpu.SUM_SEQUENTIAL)
        try:
            seq = gpu.call("extract_wave", cube, roi).download()
        finally:
            gpu.set_sum_mode(gpu.SUM_COMPENSATED)
        assert_bit_equal(seq, orc.extract_wave(host, px), f"extract_wave sequential {name} {shape}")
    # make_slice(&[(2, 0.0, 1.0)]): unit kept, only FITS axis 2 kept (at index 2), tag dropped, spectral WCS unchanged
    assert got.unit == "Jy/beam" and got.has_meta and got.tag is None
    assert got.axis(2) == ("FREQ", "Hz") and got.axis(0) == ("", "")
    assert got.pix2world(2, 3.0) == cube.pix2world(2, 3.0)


def test_roi_value_reused_across_frame_shapes(gpu, orc):
    """A Roi value keeps the device copy of its offsets (ROI::filter depends on the frame shape: roi.rs:25-34); reusing
    the value on a cube of another shape must rebuild them, and reusing it on the same shape must not change a bit."""
    px = [(3, 2), (40, 1), (7, 30), (3, 2), (63, 63), (20, 20)]
    roi = gpu.Value(gpu.Roi(px))
    for shape in [(12, 64, 64), (12, 64, 64), (9, 32, 48), (12, 64, 64)]:
        host = orc.gen_cube(0xAF1A0032, *shape)
        cube = gpu.Image.upload(host)
        gpu.set_sum_mode(gpu.SUM_SEQUENTIAL)
        try:
            seq = gpu.call("extract_wave", cube, roi).download()
        finally:
            gpu.set_sum_mode(gpu.SUM_COMPENSATED)
        assert_bit_equal(seq, orc.extract_wave(host, px), f"reused Roi value on {shape}")
        inside = [(x, y) for x, y in px if x < shape[2] and y < shape[1]]
        e = gpu.get_transform("extrude").call_results(cube, roi)[0]
        if len(inside) == len(px):
            assert_bit_equal(e.download(), orc.extrude(host, px), f"extrude, reused Roi value on {shape}")
        else:  # skipped pixels leave the vector short: from_shape_vec(..).unwrap() panics (lib.rs:1233-1256)
            assert isinstance(e, gpu.AflakError) and e.status == gpu.ERR_PANIC


def test_extract_wave_errors(gpu, orc):
    host = orc.gen_cube(0xAF1A0032, 5, 6, 7)
    cube = gpu.Image.upload(host)
    e = gpu.get_transform("extract_wave").call_results(cube, gpu.Roi())[0]  # ROI::All on a non-square frame: unwrap on None
    assert isinstance(e, gpu.AflakError) and e.status == gpu.ERR_PANIC
    e = gpu.get_transform("extract_wave").call_results(gpu.Image.upload(host[0]), gpu.Roi([(0, 0)]))[0]
    assert e.status == gpu.ERR_UNEXPECTED_INPUT and e.message == "'image' is a 2-dimensional image, while dimension 3 was expected"
    # default ROI comes from the slot (roi::ROI::All)
    sq = gpu.Image.upload(host[:, :6, :6].copy())
    assert np.allclose(gpu.get_transform("extract_wave").call_results(sq, None)[0].download(), host[:, :6, :6].sum(axis=(1, 2)), rtol=1e-5)


@pytest.mark.parametrize("shape", [(37, 64, 64), (200, 49, 49), (9, 30, 52)])
def test_extrude(gpu, orc, shape):
    nz, ny, nx = shape
    host = orc.gen_cube(0xAF1A0033, nz, ny, nx, pathology=True)
    cube = cube_with_meta(gpu, host)
    rng = np.random.default_rng(33)
    for name, px in rois(rng, ny, nx):
        t = gpu.get_transform("extrude")
        if name == "partly outside":  # a skipped pixel leaves the value vector short of (nz, datalen): unwrap panics
            e = t.call_results(cube, gpu.Roi(px))[0]
            assert isinstance(e, gpu.AflakError) and e.status == gpu.ERR_PANIC
            continue
        got = gpu.call("extrude", cube, gpu.Roi(px))
        assert got.shape == (nz, len(px))
        assert_bit_equal(got.download(), orc.extrude(host, px), f"extrude {name} {shape}")
        assert got.unit == "Jy/beam" and not got.has_meta and got.tag is None
    e = gpu.get_transform("extrude").call_results(cube, gpu.Roi())[0]  # ROI::All: datalen() == 0
    assert isinstance(e, gpu.AflakError) and e.status == gpu.ERR_PANIC


def test_make_plane3d_and_slice(gpu, orc):
    nz, ny, nx = 40, 33, 52
    host = orc.gen_cube(0xAF1A0034, nz, ny, nx)
    cube = cube_with_meta(gpu, host)
    cube.set_wcs_axis(0, 10.0, 5.0, 0.25)
    cube.set_wcs_axis(1, -3.0, 1.0, 0.5)
    cases = [
        ((7.0, 0.0, 0.0), (0.0, 0.0, 1.0), (0.0, 1.0, 0.0), nx, ny),        # a frame: defaults of dir1 / dir2
        ((0.0, 5.0, 0.0), (0.0, 0.0, 1.0), (1.0, 0.0, 0.0), nx, nz),        # position-velocity cut along x
        ((0.0, 0.0, 9.0), (0.0, 1.0, 0.0), (1.0, 0.0, 0.0), ny, nz),        # cut along y
        ((0.5, 1.25, 2.0), (0.3, 0.4, 0.5), (0.7, 0.1, 0.2), 30, 25),       # oblique plane, fractional coordinates
        ((3.0, 3.0, 3.0), (0.0, 0.0, 2.0), (0.0, 3.0, 0.0), 10, 8),         # strided, axis aligned
        ((0.0, 0.0, 0.0), (0.0, 0.0, 1.0), (0.0, 1.0, 0.0), 0, 3),          # empty mesh
    ]
    for p0, d1, d2, c1, c2 in cases:
        m = gpu.call("make_plane3d", p0, d1, d2, c1, c2)
        want_map = orc.make_plane3d(p0, d1, d2, c1, c2)
        assert m.xyz.shape == (c2, c1, 3) and np.array_equal(m.xyz, want_map)
        got = gpu.call("slice_3d_to_2d", cube, m)
        assert got.shape == (c2, c1)
        assert_bit_equal(got.download(), orc.slice_3d_to_2d(host, want_map), f"slice {p0} {d1} {d2}")
        assert got.unit == "Jy/beam" and got.tag is None
    # the frame cut keeps the two spatial axes with their WCS (make_slice with sliced_axes), the oblique one has no metadata
    frame = gpu.call("slice_3d_to_2d", cube, gpu.call("make_plane3d", (7.0, 0.0, 0.0), (0.0, 0.0, 1.0), (0.0, 1.0, 0.0), nx, ny))
    assert frame.has_meta and frame.axis(0) == ("RA---SIN", "deg") and frame.axis(1) == ("DEC--SIN", "deg") and frame.axis(2) == ("", "")
    assert frame.pix2world(0, 4.0) == cube.pix2world(0, 4.0) and frame.pix2world(1, 2.0) == cube.pix2world(1, 2.0)
    oblique = gpu.call("slice_3d_to_2d", cube, gpu.call("make_plane3d", (0.5, 1.25, 2.0), (0.3, 0.4, 0.5), (0.7, 0.1, 0.2), 30, 25))
    assert not oblique.has_meta
    # out of bounds: the first offending point in map order, formatted like Rust's `{}`
    t = gpu.get_transform("slice_3d_to_2d")
    bad = gpu.Map3d(np.array([[[0, 0, 0], [1, 2, 3]], [[40, 0.5, 2], [99, 99, 99]]], np.float32))
    e = t.call_results(cube, bad)[0]
    assert isinstance(e, gpu.AflakError) and e.status == gpu.ERR_UNEXPECTED_INPUT and e.message == "Input maps to out of bound pixel!: [40, 0.5, 2]"
    with pytest.raises(orc.OracleError) as oe:
        orc.slice_3d_to_2d(host, bad.xyz)
    assert oe.value.msg == e.message
    e = t.call_results(gpu.Image.upload(host[0]), bad)[0]
    assert e.message == "'image' is a 2-dimensional image, while dimension 3 was expected"
    e = gpu.get_transform("make_plane3d").call_results((0, 0, 0), (0, 0, 1), (0, 1, 0), -1, 3)[0]
    assert e.status == gpu.ERR_UNEXPECTED_INPUT and e.message == "'count1' must be positive, but got -1"


def test_velocity_field_map(gpu, orc):
    rng = np.random.default_rng(35)
    a = rng.uniform(6500, 6600, (61, 47)).astype(np.float32)
    a[3, 3] = np.nan
    img = gpu.Image.upload(a)
    img.unit = "Ang"
    img.tag = "BPT"
    img.set_axis(0, "RA", "deg")
    for w0 in [6562.8, 1.0, 0.0]:
        got = gpu.call("create_velocity_field_map", img, w0)
        assert_bit_equal(got.download(), orc.velocity_field(a, w0), f"velocity w0={w0}")
    assert got.unit == "km/s" and got.tag == "BPT" and got.axis(0) == ("RA", "deg") and got.homogeneous
    img.unit = "pix"
    got = gpu.call("create_velocity_field_map", img, 6562.8)
    assert got.unit == "pix*km/s/Ang" or got.unit == "km/s*pix/Ang" or "km/s" in got.unit
    assert not got.homogeneous


def test_histogram_transformation(gpu, orc):
    rng = np.random.default_rng(36)
    a = rng.lognormal(0.0, 1.0, (130, 77)).astype(np.float32)
    a[::9, ::7] = np.nan
    img = gpu.Image.upload(a)
    img.unit = "adu"
    img.tag = "color_image"
    for params in [(0.0, 0.5, 1.0), (0.1, 0.25, 0.9), (0.0, 0.03, 0.4), (0.5, 0.5, 0.5), (0.9, 0.7, 0.1)]:
        got = gpu.call("histogram_transformation", img, params, (0.0, 1.0))
        assert_bit_equal(got.download(), orc.histogram_transformation(a, params), f"histogram_transformation {params}")
    assert got.unit == "adu" and got.tag == "color_image"  # image.clone()
    rgb = rng.uniform(0, 65535, (3, 40, 52)).astype(np.float32)
    rgb[1] *= 0.5
    rgb[2, 5, 5] = np.nan
    p9 = (0.0, 0.5, 1.0, 0.1, 0.3, 0.8, 0.2, 0.7, 0.95)
    cimg = gpu.Image.upload(rgb)
    cimg.tag = "color_image"
    cimg.unit = "adu"
    got = gpu.call("histogram_transformation_rgb", cimg, p9, (0.0, 1.0))
    assert_bit_equal(got.download(), orc.histogram_transformation_rgb(rgb, p9), "histogram_transformation_rgb")
    assert got.unit == "" and got.tag == "color_image" and not got.has_meta
    e = gpu.get_transform("histogram_transformation_rgb").call_results(gpu.Image.upload(rgb[0]), p9, (0.0, 1.0))[0]
    assert isinstance(e, gpu.AflakError) and e.status == gpu.ERR_PANIC


def test_colour_image_operators(gpu, orc):
    rng = np.random.default_rng(37)
    r, g, b = (rng.uniform(0, 65535, (45, 36)).astype(np.float32) for _ in range(3))
    g[4, 4] = np.nan
    r[7, 7] = g[7, 7] = b[7, 7] = 123.0   # grey pixel: 0/0 hue
    r[8, 8] = g[8, 8] = b[8, 8] = 0.0     # black: l = 0 in the stretch, max = 0 in the saturation
    rgb = gpu.call("color_image_from_rgb", gpu.Image.upload(r), gpu.Image.upload(g), gpu.Image.upload(b))
    host = np.stack([r, g, b])
    assert rgb.shape == (3, 45, 36) and rgb.unit == "" and not rgb.has_meta
    assert_bit_equal(rgb.download(), host, "color_image_from_rgb")
    assert_bit_equal(gpu.call("color_image_to_hsv", rgb).download(), orc.color_image_to_hsv(host), "color_image_to_hsv")
    rgb.tag = "color_image"
    for beta in [1e-9, 1e-4, 0.5, 10.0]:
        got = gpu.call("apply_arcsinh_stretch", rgb, beta)
        assert_ulp(got.download(), orc.arcsinh_stretch(host, beta), 1, f"arcsinh beta={beta}")
    assert got.tag == "color_image" and got.unit == ""
    # odd pixel count -> scalar lanes; a 4-channel input keeps its extra channel untouched (out = image.clone())
    four = rng.uniform(1, 100, (4, 5, 7)).astype(np.float32)
    assert_ulp(gpu.call("apply_arcsinh_stretch", gpu.Image.upload(four), 0.1).download(), orc.arcsinh_stretch(four, 0.1), 1, "arcsinh 4 channels")
    t = gpu.get_transform("color_image_from_rgb")
    e = t.call_results(gpu.Image.upload(r), gpu.Image.upload(g[:, :-1]), gpu.Image.upload(b))[0]
    assert e.status == gpu.ERR_UNEXPECTED_INPUT and "Cannot compose images of different dimensions" in e.message
    e = t.call_results(gpu.Image.upload(host), gpu.Image.upload(g), gpu.Image.upload(b))[0]
    assert e.message == "'image_r' is a 3-dimensional image, while dimension 2 was expected"
    e = gpu.get_transform("color_image_to_hsv").call_results(gpu.Image.upload(host[:2]))[0]
    assert isinstance(e, gpu.AflakError) and e.status == gpu.ERR_PANIC


def test_waveform_template_graph(gpu, orc):
    """The reference's default "waveform" template (src/templates.rs:17-141): open_fits -> fits_to_image -> extract_wave(roi)."""
    import os
    gold = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    g = gpu.Graph()
    c_path = g.add_constant(gpu.Paths([os.path.join(gold, "test_fits_crop.fits")]))
    t_open, t_img, t_wave = g.add_transform("open_fits"), g.add_transform("fits_to_image"), g.add_transform("extract_wave")
    px = [(x, y) for y in range(3, 9) for x in range(2, 10)]
    c_roi = g.add_constant(gpu.Roi(px))
    g.connect(c_path, 0, t_open, 0)
    g.connect(t_open, 0, t_img, 0)
    g.connect(t_img, 0, t_wave, 0)
    g.connect(c_roi, 0, t_wave, 1)
    wave = g.compute(t_wave, 0)
    crop = np.load(os.path.join(gold, "test_fits_crop.npy"))
    want = orc.extract_wave(crop, px, f64=True)
    assert wave.shape == (200,) and np.allclose(wave.download(), want, rtol=1e-5, atol=1e-5)
    assert wave.unit == "K" and wave.axis(2)[0] == "VELO-LSR"
    # moving the region only recomputes extract_wave: the cube stays cached on the device
    calls = g.calls
    g.set_constant(c_roi, gpu.Roi(px[:10]))
    wave2 = g.compute(t_wave, 0)
    assert g.calls == calls + 1
    assert np.allclose(wave2.download(), orc.extract_wave(crop, px[:10], f64=True), rtol=1e-5, atol=1e-5)
