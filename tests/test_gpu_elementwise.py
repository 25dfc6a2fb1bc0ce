"""GPU parity: per-pixel image operators through the operator registry and the C ABI.

Reference: aflak_primitives/src/lib.rs:180-197 (image_min_max), 257-308 (clip/normalize/replace_nan/logscale/log10/
negation), 309-316 + 1858-1871 (apply_tone_curve), 334-359 (linear_composition, image_multiplier, slice_one_frame),
498-508 + 1584-1596 (create_equivalent_width), 554-561 + 1894-1962 (down_sampling); unit.rs:754-918 (operators).
Tier A (bit-exact vs the strict-f32 oracle): every operator built from IEEE + - * / and selects, slicing, gather.
Tier B (<= 2 ulp vs the oracle): operators that call libm (ln, log10, powf) -- CUDA's libm differs in the last bits.
"""
import numpy as np
import pytest

from util import assert_bit_equal, assert_ulp

pytestmark = pytest.mark.gpu

SHAPES = [(74, 74), (49, 49), (1, 1), (3, 1021), (256, 256), (5, 6, 7)]


def images(orc, shape, seed, pathology=False):
    n = int(np.prod(shape))
    cube = orc.gen_cube(seed, 2, 1, n, pathology=pathology)
    u, v = cube[0, 0].copy(), cube[1, 0].copy()
    if pathology:  # the generator's 0.1 % pathology rate is too thin for small images: plant every special value
        special = np.array([np.nan, np.inf, -np.inf, 0.0, -0.0, 1e-40, -3.5, np.finfo(np.float32).max], np.float32)
        rng = np.random.default_rng(seed)
        for arr in (u, v):
            idx = rng.integers(0, n, size=min(n, 16))
            arr[idx] = special[rng.integers(0, special.size, size=idx.size)]
    return u.reshape(shape), v.reshape(shape)


@pytest.mark.parametrize("shape", SHAPES)
@pytest.mark.parametrize("pathology", [False, True])
def test_linear_composition_ratio_ew_bit_exact(gpu, orc, shape, pathology):
    u, v = images(orc, shape, 11, pathology)
    du, dv = gpu.Image.upload(u), gpu.Image.upload(v)
    for a, b in [(1.0, 1.0), (0.4666667, 0.5333333), (-2.5, 1e-3), (0.0, 3.0)]:
        got = gpu.call("linear_composition", du, dv, a, b).download()
        assert_bit_equal(got, orc.linear_composition(u, v, a, b), f"linear_composition a={a} b={b}")
    assert_bit_equal(gpu.call("ratio", du, dv).download(), orc.ratio(u, v), "ratio")
    for fl, mx, em in [(20.0, 1e10, True), (1.0, float("inf"), False), (20.0, 3.0, False), (7.5, -1.0, True)]:
        got = gpu.call("create_equivalent_width", du, dv, fl, mx, em).download()
        assert_bit_equal(got, orc.equivalent_width(u, v, fl, mx, em), f"EW fl={fl} max={mx} em={em}")


def test_elementwise_kat(gpu):
    """Hand-derived answers (tests/golden/kat.json): EW(off=2,on=1,fl=20,emission) = -10; 2*1.5 + 0.25*(-2) = 2.5."""
    off = gpu.Image.upload(np.full((3, 4), 2.0, np.float32))
    on = gpu.Image.upload(np.full((3, 4), 1.0, np.float32))
    assert np.all(gpu.call("create_equivalent_width", off, on, 20.0, 1e10, True).download() == -10.0)
    assert np.all(gpu.call("create_equivalent_width", off, on, 20.0, 5.0, False).download() == 0.0)
    u = gpu.Image.upload(np.full((2, 2), 1.5, np.float32))
    v = gpu.Image.upload(np.full((2, 2), -2.0, np.float32))
    assert np.all(gpu.call("linear_composition", u, v, 2.0, 0.25).download() == 2.5)
    # IEEE division semantics: x/0 = +-Inf, 0/0 = NaN, no error
    a = gpu.Image.upload(np.array([1.0, -1.0, 0.0, 4.0], np.float32))
    b = gpu.Image.upload(np.array([0.0, 0.0, 0.0, 2.0], np.float32))
    r = gpu.call("ratio", a, b).download()
    assert r[0] == np.inf and r[1] == -np.inf and np.isnan(r[2]) and r[3] == 2.0


def test_shape_mismatch_error(gpu):
    a = gpu.Image.upload(np.zeros((2, 3), np.float32))
    b = gpu.Image.upload(np.zeros((3, 2), np.float32))
    msg = "Cannot compose images of different dimensions ('i1' has dimension [2, 3], while 'i2' has dimension [3, 2])"
    for name, args in [("linear_composition", (a, b, 1.0, 1.0)), ("ratio", (a, b)), ("image_multiplier", (a, b, 1.0, 1.0))]:
        res = gpu.get_transform(name).call_results(*args)
        assert isinstance(res[0], gpu.AflakError) and res[0].status == gpu.ERR_UNEXPECTED_INPUT
        assert res[0].message == msg, res[0].message


@pytest.mark.parametrize("shape", [(74, 74), (3, 1021), (4, 5, 6)])
def test_image_min_max(gpu, orc, shape):
    u, _ = images(orc, shape, 12)
    mn, mx = gpu.call("image_min_max", gpu.Image.upload(u))
    omn, omx = orc.image_min_max(u)
    assert (mn, mx) == (omn, omx)
    p, _ = images(orc, shape, 13, pathology=True)
    p.reshape(-1)[0] = np.nan
    mn, mx = gpu.call("image_min_max", gpu.Image.upload(p))
    omn, omx = orc.image_min_max(p)
    assert (mn == omn or (np.isnan(mn) and np.isnan(omn))) and (mx == omx)


def test_image_min_max_all_nan_and_inf(gpu):
    """Rust f32::min/max ignore NaN; the running values start at f32::MAX / f32::MIN (lib.rs:186-195)."""
    fmax = np.finfo(np.float32).max
    mn, mx = gpu.call("image_min_max", gpu.Image.upload(np.full((5, 5), np.nan, np.float32)))
    assert (mn, mx) == (fmax, -fmax)
    mn, mx = gpu.call("image_min_max", gpu.Image.upload(np.array([np.inf, 1.0, -np.inf, np.nan], np.float32)))
    assert mn == -np.inf and mx == np.inf
    mn, mx = gpu.call("image_min_max", gpu.Image.upload(np.array([-0.0, 0.0], np.float32)))
    assert mn == 0.0 and mx == 0.0


@pytest.mark.parametrize("shape", [(74, 74), (3, 1021), (2, 9, 9)])
def test_logscale_log10_multiplier_within_2ulp(gpu, orc, shape):
    u, v = images(orc, shape, 14)
    du, dv = gpu.Image.upload(u), gpu.Image.upload(v)
    vmin, vmax = orc.image_min_max(u)
    for a in (1000.0, 10.0, 2.5):
        got = gpu.call("convert_to_logscale", du, a, float(vmin), float(vmax)).download()
        want = orc.convert_to_logscale(u, a, vmin, vmax)
        # ln(a*d+1)/ln(a): 2 ulp on the logarithm; near d=0 the quotient is tiny, so compare with an absolute floor too
        close = np.abs(got.astype(np.float64) - want) <= 4 * 2.0 ** -24 * np.maximum(np.abs(want), 2.0 ** -10)
        assert close.all(), np.abs(got - want).max()
    assert_ulp(gpu.call("log10", du).download(), orc.log10(u), 2, "log10")
    for a, b in [(1.0, -1.0), (2.0, 0.5), (0.5, 1.5)]:
        assert_ulp(gpu.call("image_multiplier", du, dv, a, b).download(), orc.image_multiplier(u, v, a, b), 4, f"multiplier {a},{b}")
    # the registry spelling of `ratio`: u^1 * v^-1 equals u/v to within powf accuracy
    assert_ulp(gpu.call("image_multiplier", du, dv, 1.0, -1.0).download(), orc.ratio(u, v), 4, "u^1 v^-1 vs u/v")
    # exponents with a closed form (1 = the default of both slots, -1, 2, 0) skip powf on the device: same answers as libm's powf
    # on every special value (C99 F.9.4.4: pow(x, 0) = 1 even for NaN, pow(+-0, -1) = +-Inf, pow(-Inf, 2) = +Inf ...)
    p, q = images(orc, shape, 16, pathology=True)
    dp, dq = gpu.Image.upload(p), gpu.Image.upload(q)
    for a, b in [(1.0, 1.0), (2.0, 0.0), (0.0, -1.0), (-1.0, 2.0), (2.0, 2.0), (1.0, 0.5), (-1.0, -1.0)]:
        for x, y, dx, dy in ((u, v, du, dv), (p, q, dp, dq)):
            assert_ulp(gpu.call("image_multiplier", dx, dy, a, b).download(), orc.image_multiplier(x, y, a, b), 4, f"multiplier {a},{b}")
    with np.errstate(all="ignore"):  # Inf * 0 and overflow are part of the case
        plain = (p * q).astype(np.float32)
    assert_bit_equal(gpu.call("image_multiplier", dp, dq, 1.0, 1.0).download(), plain, "u^1 v^1 is the plain product")


def test_logscale_kat_and_nonfinite(gpu):
    img = gpu.Image.upload(np.array([2.0, 9.0, np.nan, 1.0], np.float32))
    out = gpu.call("convert_to_logscale", img, 1000.0, 2.0, 9.0).download()
    assert out[0] == 0.0                                      # v = vmin -> ln(1)/ln(a) = 0
    assert abs(out[1] - np.log(1001.0) / np.log(1000.0)) < 1e-6
    assert np.isnan(out[2]) and np.isnan(out[3])              # ln of a negative argument
    out = gpu.call("log10", gpu.Image.upload(np.array([100.0, 0.0, -1.0, np.inf], np.float32))).download()
    assert out[0] == 2.0 and out[1] == -np.inf and np.isnan(out[2]) and out[3] == np.inf


@pytest.mark.parametrize("shape", [(74, 74), (3, 1021)])
def test_apply_tone_curve_bit_exact(gpu, orc, shape):
    u, _ = images(orc, shape, 15)
    img = (u * 20000.0).astype(np.float32)
    img.reshape(-1)[:4] = [0.0, 65535.0, 257.0, -3.0]
    d = gpu.Image.upload(img)
    assert_bit_equal(gpu.call("apply_tone_curve", d, gpu.ToneCurve()).download(), orc.apply_tone_curve(img, orc.tone_curve_default()))
    for cp in [[(0.0, 0.0), (0.5, 0.8), (1.0, 1.0)], [(0.0, 0.1), (1.0, 0.9)], [(0.0, 0.0), (0.3, 0.1), (0.6, 0.9), (1.0, 1.0)]]:
        table = orc.create_curve(cp)
        got = gpu.call("apply_tone_curve", d, gpu.ToneCurve(table, cp)).download()
        assert_bit_equal(got, orc.apply_tone_curve(img, table), f"tone curve {cp}")
    with_nan = img.copy()
    with_nan.reshape(-1)[5] = np.nan  # NaN as usize == 0 -> table[0]
    got = gpu.call("apply_tone_curve", gpu.Image.upload(with_nan), gpu.ToneCurve()).download()
    assert_bit_equal(got, orc.apply_tone_curve(with_nan, orc.tone_curve_default()))


def test_apply_tone_curve_out_of_range_is_an_error_not_ub(gpu, orc):
    """The reference indexes out of bounds and panics for v >~ 65535*(1+1/255); the ABI reports it."""
    img = np.array([10.0, 70000.0], np.float32)
    res = gpu.get_transform("apply_tone_curve").call_results(gpu.Image.upload(img), gpu.ToneCurve())
    assert isinstance(res[0], gpu.AflakError) and res[0].status == gpu.ERR_PANIC
    with pytest.raises(orc.OracleError) as ei:
        orc.apply_tone_curve(img, orc.tone_curve_default())
    assert ei.value.code == orc.ERR_PANIC


def test_slice_one_frame_is_a_view_and_bit_exact(gpu, orc):
    host = orc.gen_cube(16, 9, 5, 8, pathology=True)
    cube = gpu.Image.upload(host)
    for f in (0, 4, 8):
        fr = gpu.call("slice_one_frame", cube, f)
        assert fr.shape == (5, 8)
        assert_bit_equal(fr.download(), orc.slice_one_frame(host, f))
        assert fr.device_ptr == cube.device_ptr + f * 5 * 8 * 4  # zero-copy: aliases the cube
    fr = gpu.call("slice_one_frame", cube, 3)
    del cube  # the view keeps its parent alive
    assert_bit_equal(fr.download(), host[3])
    cube = gpu.Image.upload(host)
    t = gpu.get_transform("slice_one_frame")
    assert t.call_results(cube, 9)[0].message == "'frame_idx' greater or equal to the input image 'input_img''s frame count (expected 9 < 9)"
    assert t.call_results(cube, -1)[0].message == "'frame_idx' must be positive, but got -1"
    # a 2-D image slices to a 1-D row
    assert gpu.call("slice_one_frame", gpu.Image.upload(host[0]), 2).shape == (8,)


def test_linear_composition_of_three_slices(gpu, orc):
    """BASELINE config 2, second half: linear_composition of 3 slice_one_frame's = two chained calls."""
    host = orc.gen_cube(17, 200, 74, 74)
    cube = gpu.Image.upload(host)
    f = [gpu.call("slice_one_frame", cube, k) for k in (109, 144, 184)]
    t = gpu.call("linear_composition", f[0], f[1], 0.25, 0.5)
    out = gpu.call("linear_composition", t, f[2], 1.0, 0.25).download()
    want = orc.linear_composition(orc.linear_composition(host[109], host[144], 0.25, 0.5), host[184], 1.0, 0.25)
    assert_bit_equal(out, want)


@pytest.mark.parametrize("shape", [(4, 5), (74, 74), (49, 49), (1, 7), (3, 10, 11), (2, 64, 64)])
@pytest.mark.parametrize("n", [1, 2, 3, 7, 100])
def test_down_sampling_bit_exact(gpu, orc, shape, n):
    u, _ = images(orc, shape, 18, pathology=True)
    d = gpu.Image.upload(u)
    d.tag = "color_image"
    d.unit = "K"
    out = gpu.call("down_sampling", d, n)
    want = orc.down_sampling(u, n)
    assert out.shape == want.shape
    assert_bit_equal(out.download(), want, f"down_sampling {shape} n={n}")
    assert out.tag == "color_image" and out.unit == "" and not out.has_meta  # tag kept, unit and WCS dropped


def test_down_sampling_errors(gpu):
    t = gpu.get_transform("down_sampling")
    img = gpu.Image.upload(np.zeros((4, 4), np.float32))
    assert t.call_results(img, 0)[0].message == "n should be greater than 0, 0 > 0"
    assert t.call_results(img, -2)[0].message == "n should be greater than 0, -2 > 0"
    assert t.call_results(gpu.Image.upload(np.zeros(4, np.float32)), 2)[0].status == gpu.ERR_UNEXPECTED_INPUT


def test_adjacent_elementwise_operators(gpu):
    """negation / normalize / clip / replace_nan (lib.rs:257-308): IEEE-only arithmetic -> exact against numpy f32."""
    rng = np.random.default_rng(3)
    x = rng.standard_normal((37, 41)).astype(np.float32)
    x[0, :3] = [np.nan, np.inf, -np.inf]
    d = gpu.Image.upload(x)
    assert_bit_equal(gpu.call("negation", d).download(), -x)
    lo, hi = np.float32(-0.5), np.float32(2.0)
    assert_bit_equal(gpu.call("normalize_image", d, float(lo), float(hi)).download(), (x - lo) / (hi - lo))
    assert_bit_equal(gpu.call("replace_nan_image", d, 7.0).download(), np.where(np.isnan(x), np.float32(7.0), x))
    got = gpu.call("clip_image", d, 1.0, True, -1.0, True).download()
    want = np.where((x >= 1.0) | (x <= -1.0), np.float32(np.nan), x)
    assert_bit_equal(got, want)
    got = gpu.call("clip_image", d, 1.0, True, -1.0, False).download()
    assert_bit_equal(got, np.where(x >= 1.0, np.float32(np.nan), x))
    # range_specification: sub-cube copy
    cube = rng.standard_normal((10, 3, 5)).astype(np.float32)
    out = gpu.call("range_specification", gpu.Image.upload(cube), 2, 7)
    assert out.shape == (5, 3, 5)
    assert_bit_equal(out.download(), cube[2:7])


def test_metadata_rules(gpu, orc):
    """Host-side metadata travels with the handle exactly as the reference's operators define it."""
    u, v = images(orc, (6, 6), 19)
    du, dv = gpu.Image.upload(u), gpu.Image.upload(v)
    for d in (du, dv):
        d.unit = "K"
        d.tag = "BPT"
        d.set_axis(0, "RA", "deg")
        d.set_axis(1, "DEC", "deg")
    out = gpu.call("linear_composition", du, dv, 1.0, 1.0)
    assert out.unit == "K" and out.tag is None and out.has_meta and out.homogeneous
    dv.unit = "Jy"
    assert not gpu.call("linear_composition", du, dv, 1.0, 1.0).homogeneous     # unit.rs:659-673
    dv.set_axis(0, "GLON", "deg")
    assert not gpu.call("linear_composition", du, dv, 1.0, 1.0).has_meta        # metas differ -> None (unit.rs:817-832)
    r = gpu.call("ratio", du, dv)
    assert r.unit == "K/Jy" and r.has_meta and r.tag is None                     # unit.rs:863-880
    du2 = gpu.Image.upload(u)
    du2.unit = "K"
    assert gpu.call("ratio", du, du2).unit == ""                                 # K/K cancels
    ew = gpu.call("create_equivalent_width", du, du2, 1.0, 1e10, False)
    assert ew.unit == "Ang" and not ew.has_meta and ew.tag is None              # K*Ang/K
    ls = gpu.call("convert_to_logscale", du, 1000.0, 0.0, 1.0)
    assert ls.unit == "K" and ls.tag == "BPT" and ls.has_meta                    # clone keeps everything
    tc = gpu.call("apply_tone_curve", du, gpu.ToneCurve())
    assert tc.unit == "" and tc.tag is None and not tc.has_meta                 # from_array(.., Unit::None)
    ct = gpu.call("change_tag", du, "color_image")
    assert ct.tag == "color_image" and ct.unit == "K" and ct.device_ptr == du.device_ptr


def test_signed_zeros_are_results_too(gpu, orc):
    """-0.0 and +0.0 are different bit patterns and the reference's IEEE arithmetic distinguishes them (0 * -1 = -0,
    -0 + 0 = +0, -0 / x = -0): assert_bit_equal compares raw bits, and every tier-A operator is fed the zero cases."""
    from util import bits

    z = np.array([0.0, -0.0, 0.0, -0.0, 1.0, -1.0, 0.0, -0.0], np.float32)
    w = np.array([0.0, 0.0, -0.0, -0.0, 0.0, -0.0, 2.0, -3.0], np.float32)
    dz, dw = gpu.Image.upload(z), gpu.Image.upload(w)
    for a, b in [(1.0, 1.0), (-1.0, 1.0), (0.0, 0.0), (-0.0, 2.0), (3.0, -0.0)]:
        assert_bit_equal(gpu.call("linear_composition", dz, dw, a, b).download(), orc.linear_composition(z, w, a, b), f"lincomp {a} {b}")
    assert_bit_equal(gpu.call("ratio", dz, dw).download(), orc.ratio(z, w), "ratio")
    for fl, mx, em in [(20.0, 1e10, True), (20.0, 1e10, False), (0.0, -1.0, True), (-0.0, 0.0, False)]:
        assert_bit_equal(gpu.call("create_equivalent_width", dz, dw, fl, mx, em).download(), orc.equivalent_width(z, w, fl, mx, em), f"EW {fl} {mx} {em}")
    neg = gpu.call("negation", dz).download()
    assert np.array_equal(bits(neg), bits(z) ^ np.uint32(0x80000000))
    assert_bit_equal(gpu.call("replace_nan_image", dz, -0.0).download(), z)
    assert_bit_equal(gpu.call("replace_nan_image", gpu.Image.upload(np.array([np.nan, -0.0], np.float32)), -0.0).download(),
                     np.array([-0.0, -0.0], np.float32))
    assert_bit_equal(gpu.call("normalize_image", dz, 0.0, 1.0).download(), ((z - np.float32(0.0)) / np.float32(1.0)).astype(np.float32))
    assert_bit_equal(gpu.call("normalize_image", dz, 0.0, -2.0).download(), ((z - np.float32(0.0)) / np.float32(-2.0)).astype(np.float32))
    # copies: slice_one_frame and down_sampling move bit patterns
    cube = np.stack([z.reshape(2, 4), w.reshape(2, 4)])
    dc = gpu.Image.upload(cube)
    assert_bit_equal(gpu.call("slice_one_frame", dc, 1).download(), cube[1])
    assert_bit_equal(gpu.call("down_sampling", dc, 2).download(), cube[:, ::2, ::2])
    # the sum of frames that are all -0.0 is -0.0 (ndarray sum_axis starts from the first frame, not from +0.0)
    mz = np.full((5, 3, 4), -0.0, np.float32)
    mz[:, 0, 0] = [0.0, -0.0, -0.0, -0.0, -0.0]
    mz[:, 0, 1] = [-0.0, -0.0, 0.0, -0.0, -0.0]
    dm = gpu.Image.upload(mz)
    for mode in (gpu.SUM_SEQUENTIAL, gpu.SUM_COMPENSATED):
        gpu.set_sum_mode(mode)
        try:
            assert_bit_equal(gpu.call("integral", dm, 0, 5)[0].download(), orc.integral(mz, 0, 5), f"integral of zeros mode {mode}")
            assert_bit_equal(gpu.call("average", dm, 1, 5)[0].download(), orc.average(mz, 1, 5), f"average of zeros mode {mode}")
        finally:
            gpu.set_sum_mode(gpu.SUM_INHERIT)
