"""GPU parity: axis-0 reductions (integral / average) through the operator registry and the C ABI.

Reference: aflak_primitives/src/lib.rs:384-411, 1303-1330; precond.rs:3-98.
Tier C tolerance: |gpu - o64| <= 1e-5*|o64| or <= 4*2^-24*sum|x|, stated in tests/util.py.
In SUM_SEQUENTIAL mode the result must be bit-identical to the oracle's left-to-right f32 sum.
"""
import numpy as np
import pytest

from util import assert_bit_equal, assert_reduction

pytestmark = pytest.mark.gpu

SHAPES = [
    (64, 8, 8),        # tiny
    (200, 49, 49),     # the reference's test.fits shape; 49*49 is odd -> scalar lanes
    (131, 37, 52),     # n2 % 4 == 0, ragged tile
    (457, 74, 74),     # MaNGA spatial size, split-K
    (33, 1, 1),        # a single spaxel
    (7, 300, 301),     # wide, shallow
]


@pytest.mark.parametrize("shape", SHAPES)
@pytest.mark.parametrize("name", ["integral", "average"])
def test_reduction_compensated(gpu, orc, shape, name):
    nz, ny, nx = shape
    host = orc.gen_cube(0xAF1A0001, nz, ny, nx)
    cube = gpu.Image.upload(host)
    for start, end in [(0, nz), (0, 1), (nz - 1, nz), (nz // 3, 2 * nz // 3 + 1), (1, min(nz, 10))]:
        if not start < end:
            continue
        img, middle, width = gpu.call(name, cube, start, end)
        s64, a64 = orc.sum_f64(host, start, end)
        div = float(end - start) if name == "average" else 1.0
        assert_reduction(img.download(), s64, a64, f"{name}{shape}[{start},{end})", divisor=div)
        assert (float(middle), float(width)) == orc.middle_width(start, end)
        assert img.shape == (ny, nx)


@pytest.mark.parametrize("shape", SHAPES)
@pytest.mark.parametrize("name", ["integral", "average"])
def test_reduction_sequential_bit_exact(gpu, orc, shape, name):
    nz, ny, nx = shape
    host = orc.gen_cube(0xAF1A0002, nz, ny, nx)
    cube = gpu.Image.upload(host)
    gpu.set_sum_mode(gpu.SUM_SEQUENTIAL)
    try:
        for start, end in [(0, nz), (nz // 2, nz), (0, 1)]:
            got = gpu.call(name, cube, start, end)[0].download()
            want = getattr(orc, name)(host, start, end)
            assert_bit_equal(got, want, f"{name}{shape}[{start},{end})")
    finally:
        gpu.set_sum_mode(gpu.SUM_COMPENSATED)


@pytest.mark.parametrize("mode", [0, 1])
def test_reduction_pathology_nan_inf_positions(gpu, orc, mode):
    """NaN / +-Inf / zeros / negatives / denormals: non-finite positions must match the reference's f32 sum."""
    nz, ny, nx = 300, 40, 44
    host = orc.gen_cube(0xAF1A0003, nz, ny, nx, pathology=True)
    assert np.isnan(host).any() and np.isinf(host).any()
    cube = gpu.Image.upload(host)
    gpu.set_sum_mode(mode)
    try:
        for start, end in [(0, nz), (10, 40), (100, 101)]:
            got = gpu.call("integral", cube, start, end)[0].download()
            want = orc.integral(host, start, end)
            assert np.array_equal(np.isnan(got), np.isnan(want))
            inf = np.isinf(want)
            assert np.array_equal(got[inf], want[inf])
            if mode == 1:
                assert_bit_equal(got, want, "pathology sequential")
            else:
                fin = np.isfinite(want)
                s64, a64 = orc.sum_f64(host, start, end)
                err = np.abs(got[fin].astype(np.float64) - s64[fin])
                assert np.all((err <= 1e-5 * np.abs(s64[fin])) | (err <= 4 * 2.0 ** -24 * a64[fin]))
    finally:
        gpu.set_sum_mode(gpu.SUM_COMPENSATED)


def test_known_answers(gpu):
    """Hand-derivable KATs: cube[k,y,x] = k -> integral over [a,b) = (a+b-1)(b-a)/2 exactly."""
    nz, ny, nx = 64, 5, 12
    host = np.broadcast_to(np.arange(nz, dtype=np.float32)[:, None, None], (nz, ny, nx)).copy()
    cube = gpu.Image.upload(host)
    for a, b in [(0, 64), (3, 17), (10, 11)]:
        got = gpu.call("integral", cube, a, b)[0].download()
        assert np.all(got == np.float32((a + b - 1) * (b - a) / 2))
        avg = gpu.call("average", cube, a, b)[0].download()
        assert np.all(avg == np.float32(np.float32((a + b - 1) * (b - a) / 2) / np.float32(b - a)))
    # average over [a, a+1) is frame a bit-exact
    rnd = np.random.default_rng(1).standard_normal((9, 6, 7)).astype(np.float32)
    c2 = gpu.Image.upload(rnd)
    assert_bit_equal(gpu.call("average", c2, 4, 5)[0].download(), rnd[4])


def test_lower_and_higher_rank(gpu, orc):
    """reduce_array_slice works for any ndim >= 1 (reduces axis 0)."""
    rng = np.random.default_rng(2)
    v = rng.standard_normal(50).astype(np.float32)
    out = gpu.call("integral", gpu.Image.upload(v), 5, 45)[0]
    assert out.shape == ()
    assert abs(float(out.download()) - float(v[5:45].astype(np.float64).sum())) < 1e-4
    m = rng.standard_normal((20, 9)).astype(np.float32)
    out = gpu.call("integral", gpu.Image.upload(m), 0, 20)[0]
    assert out.shape == (9,)
    np.testing.assert_allclose(out.download(), m.astype(np.float64).sum(0), rtol=1e-5, atol=1e-6)
    h = rng.standard_normal((6, 3, 4, 8)).astype(np.float32)
    out = gpu.call("average", gpu.Image.upload(h), 1, 5)[0]
    assert out.shape == (3, 4, 8)
    np.testing.assert_allclose(out.download(), h[1:5].astype(np.float64).mean(0), rtol=1e-5, atol=1e-6)


def test_precondition_errors(gpu, orc):
    """precond.rs messages come back verbatim as IOErr::UnexpectedInput on the Image output; the two Float outputs
    are computed before the checks and stay Ok (lib.rs:393-396)."""
    host = orc.gen_cube(1, 10, 4, 4)
    cube = gpu.Image.upload(host)
    t = gpu.get_transform("integral")
    cases = [
        ((-1, 5), "'start' must be positive, but got -1"),
        ((2, -3), "'end' must be positive, but got -3"),
        ((5, 3), "Expected start < end, but got 5 < 3"),
        ((4, 4), "Expected start < end, but got 4 < 4"),
        ((0, 11), "'end' greater or equal to the input image 'image''s frame count (expected 11 < 11)"),
    ]
    for (s, e), msg in cases:
        res = t.call_results(cube, s, e)
        assert len(res) == 3
        r = res[0]
        assert isinstance(r, gpu.AflakError), (s, e, r)
        assert r.status == gpu.ERR_UNEXPECTED_INPUT
        assert r.message == msg, r.message
        assert (float(res[1]), float(res[2])) == orc.middle_width(s, e)
        with pytest.raises(orc.OracleError) as ei:
            orc.integral(host, s, e)
        assert ei.value.msg == msg
    # end == nz is legal (half-open range)
    assert gpu.call("integral", cube, 0, 10)[0].shape == (4, 4)


def test_meta_make_slice(gpu, orc):
    """integral/average -> make_slice: unit kept, tag dropped, spectral axis dropped (unit.rs:424-451)."""
    host = orc.gen_cube(1, 10, 4, 6)
    cube = gpu.Image.upload(host)
    cube.unit = "K"
    cube.tag = "BPT"
    cube.set_axis(0, "RA", "deg")
    cube.set_axis(1, "DEC", "deg")
    cube.set_axis(2, "WAVE", "Angstrom")
    out = gpu.call("average", cube, 2, 8)[0]
    assert out.unit == "K"
    assert out.tag is None
    assert out.has_meta
    assert out.axis(0) == ("RA", "deg") and out.axis(1) == ("DEC", "deg")


def test_float_arguments_are_converted(gpu, orc):
    """cake feed converts Float -> Integer with round() (aflak_primitives/src/lib.rs:795-823)."""
    host = orc.gen_cube(1, 10, 4, 4)
    cube = gpu.Image.upload(host)
    a = gpu.call("integral", cube, 1.6, 7.4)[0].download()  # -> [2, 7)
    b = gpu.call("integral", cube, 2, 7)[0].download()
    assert_bit_equal(a, b)


def test_range_memo_left_by_the_reduce_kernel(gpu, orc):
    """The kernel that writes a reduction's pixels also leaves image_min_max of them beside the image (the range memo): a
    following image_min_max / get_vmin / get_vmax launches no kernel and returns the reference's values (lib.rs:180-197,
    lims.rs:7-35), the fused logscale tail takes one launch; NaN results are ignored, an all-NaN result keeps the
    operators' initial values.  Covers the stream-K kernel, its small-image variants and the cluster variant."""
    for shape, win in [((96, 37, 52), (10, 90)), ((300, 74, 74), (0, 300)), ((64, 300, 260), (3, 61)), ((40, 8, 4), (0, 40))]:
        nz, ny, nx = shape
        host = orc.gen_cube(0xAF1A0070 + nz, nz, ny, nx, pathology=True)
        cube = gpu.Image.upload(host)
        for mode in (gpu.SUM_COMPENSATED, gpu.SUM_SEQUENTIAL):
            gpu.set_sum_mode(mode)
            try:
                for op in ("integral", "average"):
                    img = gpu.call(op, cube, *win)[0]
                    got = img.download()
                    l0 = gpu.launch_count()
                    mn, mx = gpu.call("image_min_max", img)
                    vmin, vmax = img.vmin_vmax()
                    assert gpu.launch_count() == l0, "image_min_max of a reduction result must come from the range memo"
                    omn, omx = orc.image_min_max(got)
                    assert (np.float32(mn), np.float32(mx)) == (np.float32(omn), np.float32(omx)), (shape, mode, op)
                    assert np.float32(vmin) == np.float32(orc.get_vmin(got)) and np.float32(vmax) == np.float32(orc.get_vmax(got))
                    l0 = gpu.launch_count()
                    ls = gpu.fused_logscale_minmax(img, 1000.0).download()
                    assert gpu.launch_count() == l0 + 1
                    want = gpu.call("convert_to_logscale", img, 1000.0, float(mn), float(mx)).download()
                    assert np.array_equal(np.isnan(ls), np.isnan(want)) and np.array_equal(ls[~np.isnan(ls)], want[~np.isnan(want)])
            finally:
                gpu.set_sum_mode(gpu.SUM_COMPENSATED)
    # all-NaN result
    nanc = gpu.Image.upload(np.full((5, 6, 8), np.nan, np.float32))
    img = gpu.call("integral", nanc, 0, 5)[0]
    l0 = gpu.launch_count()
    mn, mx = gpu.call("image_min_max", img)
    assert gpu.launch_count() == l0
    omn, omx = orc.image_min_max(img.download())
    assert (np.float32(mn), np.float32(mx)) == (np.float32(omn), np.float32(omx))
