"""GPU parity: the aflak_plot colour-mapping path (f32 image -> RGB8 / RGBA8 bytes), byte-exact.

Reference: aflak_plot/src/imshow/lut.rs:132-194 (color_at_bounds, color_at, mtf, table build), image.rs:20-84
(make_raw_image, make_raw_image_rgb), image.rs:123-129 (vertical flip), lims.rs:7-35,108-124 (get_vmin/get_vmax),
util.rs:3-14 (clamp).  Every output byte must equal the oracle's (integer/byte work: bit-exact).
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def lognormal(shape, seed, nan_frac=0.0):
    rng = np.random.default_rng(seed)
    img = np.exp(rng.standard_normal(shape)).astype(np.float32)
    if nan_frac:
        img[rng.random(shape) < nan_frac] = np.nan
    return img


@pytest.mark.parametrize("shape", [(64, 64), (37, 53), (1, 1), (5, 1024), (300, 7)])
@pytest.mark.parametrize("gradient", ["Flame", "Grey", "HeatMap_Inv", "GreyClip"])
@pytest.mark.parametrize("channels", [3, 4])
def test_make_raw_image_bytes(gpu, orc, shape, gradient, channels):
    img = lognormal(shape, 21, nan_frac=0.01)
    d = gpu.Image.upload(img)
    lut = gpu.Lut(gradient)
    table = orc.lut_build(gradient)
    vmin, vmax = orc.get_vmin(img), orc.get_vmax(img)
    got = gpu.colormap(d, lut, channels=channels)  # vmin/vmax computed on the device (Image::new, image.rs:143-194)
    want = orc.make_raw_image(img, vmin, vmax, table, channels=channels)
    assert np.array_equal(got, want), f"{(got != want).sum()} bytes differ"
    gmn, gmx = d.vmin_vmax()
    assert (gmn, gmx) == (vmin, vmax)


# (0, 0.5, 1), (l0, 0.5, l2) and (0, l1, 1) take division-free shortcuts in lut_index: same bytes required, denormals included
@pytest.mark.parametrize("lims", [(0.0, 0.5, 1.0), (0.2, 0.3, 0.9), (0.0, 0.1, 0.5), (0.4, 0.9, 0.6), (0.2, 0.5, 0.9), (0.0, 0.3, 1.0)])
def test_lims_and_midtone_transfer(gpu, orc, lims):
    img = lognormal((96, 100), 22, nan_frac=0.001)
    img[0, :4] = [np.inf, -np.inf, 0.0, -1.0]
    img[1, :6] = [1e-40, 3e-45, -0.0, 1.1754944e-38, 1.5e-5, 0.99999994]
    d = gpu.Image.upload(img)
    lut = gpu.Lut("Flame")
    lut.set_lims(*lims)
    eff = lut.lims()  # set_min/set_max clamp against each other (lut.rs:202-236)
    table = orc.lut_build("Flame")  # built with the default lims
    for vmin, vmax in [(0.1, 5.0), (1.0, 1.0), (2.0, 0.5), (0.0, 1.0), (0.0, 1e-38)]:
        got = gpu.colormap(d, lut, vmin=vmin, vmax=vmax, channels=3)
        want = orc.make_raw_image(img, vmin, vmax, table, lims=eff, channels=3)
        assert np.array_equal(got, want), (lims, vmin, vmax, int((got != want).sum()))


def test_flip_and_nan_colour(gpu, orc):
    img = np.array([[0.0, 1.0], [np.nan, 0.5]], np.float32)
    d = gpu.Image.upload(img)
    lut = gpu.Lut("Grey")
    out = gpu.colormap(d, lut, vmin=0.0, vmax=1.0, channels=3)
    # row 0 of the output is the last input row (image.rs:128); NaN -> index 0 -> lut[0]
    assert out[0, 0].tolist() == [0, 0, 0] and out[0, 1].tolist() == [127, 127, 127]
    assert out[1, 0].tolist() == [0, 0, 0] and out[1, 1].tolist() == [255, 255, 255]
    noflip = gpu.colormap(d, lut, vmin=0.0, vmax=1.0, channels=3, flip_y=False)
    assert np.array_equal(noflip, out[::-1])
    rgba = gpu.colormap(d, lut, vmin=0.0, vmax=1.0, channels=4)
    assert np.all(rgba[..., 3] == 255) and np.array_equal(rgba[..., :3], out)


def test_all_nan_image_is_an_error(gpu):
    d = gpu.Image.upload(np.full((4, 4), np.nan, np.float32))
    mn, mx = d.vmin_vmax()
    assert np.isnan(mn) and np.isnan(mx)  # lims.rs:108-124
    with pytest.raises(gpu.AflakError) as ei:
        gpu.colormap(d, gpu.Lut("Grey"))
    assert "vmin, vmax not set" in ei.value.message  # image.rs:46-48


def test_custom_gradient_golden_through_the_kernel(gpu):
    """The reference's golden colours (lut.rs:292-343) reproduced by the device kernel, one pixel each."""
    lut = gpu.Lut([(0.0, (0, 0, 255)), (0.5, (255, 255, 255)), (1.0, (255, 0, 0))])
    pts = np.array([[0.0, 1.0, 0.5, 0.25]], np.float32)
    out = gpu.colormap(gpu.Image.upload(pts), lut, vmin=0.0, vmax=1.0, channels=3, flip_y=False)
    assert out[0].tolist() == [[0, 0, 255], [255, 0, 0], [254, 254, 255], [127, 127, 255]]
    lut2 = gpu.Lut([(0.0, (0, 0, 0)), (1.0, (255, 255, 255))])
    lut2.set_lims(0.2, 0.5, 0.9)
    pts = np.array([[0.0, 0.1, 0.2, 0.55, 0.9, 0.95, 1.0]], np.float32)
    out = gpu.colormap(gpu.Image.upload(pts), lut2, vmin=0.0, vmax=1.0, channels=3, flip_y=False)
    assert out[0, :, 0].tolist() == [0, 0, 0, 127, 255, 255, 255]


def test_make_raw_image_rgb(gpu, orc):
    """3-channel image, one LUT and one (vmin, vmax) per channel (image.rs:51-84)."""
    rng = np.random.default_rng(23)
    img = rng.random((3, 45, 52)).astype(np.float32)
    img[1, 3, 3] = np.nan
    names = ["Red", "Green", "Blue"]
    luts = [gpu.Lut(n) for n in names]
    luts[1].set_lims(0.1, 0.4, 0.8)
    tables = np.stack([orc.lut_build(n) for n in names])
    lims3 = [l.lims() for l in luts]
    vmin3, vmax3 = [0.0, 0.1, 0.2], [1.0, 0.9, 0.7]
    for ch in (3, 4):
        got = gpu.colormap_rgb(gpu.Image.upload(img), luts, vmin3, vmax3, channels=ch)
        want = orc.make_raw_image_rgb(img, vmin3, vmax3, tables, lims3, channels=ch)
        assert np.array_equal(got, want), int((got != want).sum())


def test_fused_chain_matches_unfused_operators(gpu, orc):
    """BASELINE config 4 chain: logscale -> rescale to 0..65535 -> tone curve -> colour LUT, fused into the colour-map
    kernel, against the same chain run operator by operator on the device and against the oracle."""
    img = lognormal((128, 160), 24)
    d = gpu.Image.upload(img)
    vmin, vmax = gpu.call("image_min_max", d)
    a = 1000.0
    tone = orc.create_curve([(0.0, 0.0), (0.4, 0.7), (1.0, 1.0)])
    lut = gpu.Lut("Flame")
    lut.set_lims(0.05, 0.4, 0.95)
    # unfused on the device: three operators materialise f32 images, then the plain colour map
    ls = gpu.call("convert_to_logscale", d, a, float(vmin), float(vmax))
    lmax = np.float32(np.log(np.float32(a) + np.float32(1.0)) / np.log(np.float32(a)))
    nmin, nmax = 0.0, float(lmax / np.float32(65535.0))  # normalize_image so that the result spans 0..65535
    nm = gpu.call("normalize_image", ls, nmin, nmax)
    tc = gpu.call("apply_tone_curve", nm, gpu.ToneCurve(tone))
    unfused = gpu.colormap(tc, lut, vmin=0.0, vmax=65535.0, channels=4)
    chain = {"logscale_a": a, "logscale_vmin": float(vmin), "logscale_vmax": float(vmax), "normalize_min": nmin,
             "normalize_max": nmax, "tone_table": tone}
    fused = gpu.colormap(d, lut, vmin=0.0, vmax=65535.0, channels=4, chain=chain)
    assert np.array_equal(fused, unfused), int((fused != unfused).sum())
    # logscale limits left to the library (NaN, NaN): image_min_max runs on the device and feeds the chain without a read-back
    auto = gpu.colormap(d, lut, vmin=0.0, vmax=65535.0, channels=4, chain=dict(chain, logscale_vmin=float("nan"), logscale_vmax=float("nan")))
    assert np.array_equal(auto, unfused), int((auto != unfused).sum())
    # oracle: same chain on the CPU; libm's logf may differ from CUDA's in the last bit, which can move a pixel by one
    # table entry -> allow a channel difference of at most 1 on at most 0.1 % of the pixels
    o = orc.convert_to_logscale(img, a, vmin, vmax)
    o = ((o - np.float32(nmin)) / (np.float32(nmax) - np.float32(nmin))).astype(np.float32)
    o = orc.apply_tone_curve(o, tone)
    want = orc.make_raw_image(o, 0.0, 65535.0, orc.lut_build("Flame"), lims=lut.lims(), channels=4)
    diff = np.abs(fused.astype(np.int32) - want.astype(np.int32))
    assert diff.max() <= 2 and (diff.max(axis=2) > 0).mean() < 1e-3, (diff.max(), (diff.max(axis=2) > 0).mean())


def test_fused_chain_tone_index_out_of_bounds(gpu, orc):
    """apply_tone_curve indexes its table unchecked (lib.rs:1866): an input above 65535*(1+1/(len-1)) panics in the
    reference; the fused chain reports the same error as the stand-alone operator instead of clamping silently."""
    img = lognormal((64, 64), 26)
    d = gpu.Image.upload(img)
    vmin, vmax = gpu.call("image_min_max", d)
    tone = orc.tone_curve_default()
    lut = gpu.Lut("Flame")
    # normalise so that the brightest pixel lands at twice the table range
    lmax = np.float32(np.log(np.float32(1001.0)) / np.log(np.float32(1000.0)))
    chain = {"logscale_a": 1000.0, "logscale_vmin": float(vmin), "logscale_vmax": float(vmax), "normalize_min": 0.0,
             "normalize_max": float(lmax / np.float32(2 * 65535.0)), "tone_table": tone}
    with pytest.raises(gpu.AflakError) as ei:
        gpu.colormap(d, lut, vmin=0.0, vmax=65535.0, channels=4, chain=chain)
    assert ei.value.status == gpu.ERR_PANIC and "index out of bounds" in str(ei.value)
    # the device-buffer form is stream-ordered (no read-back on the steady-state path): the panic surfaces at the next synchronise
    out = gpu.Image.alloc((64, 64))
    gpu.colormap_device(d, lut, 0.0, 65535.0, out.device_ptr, channels=4, chain=chain)
    with pytest.raises(gpu.AflakError) as ei:
        gpu.synchronize()
    assert ei.value.status == gpu.ERR_PANIC and "index out of bounds" in str(ei.value)
    gpu.synchronize()  # reported once
    nm = gpu.call("normalize_image", gpu.call("convert_to_logscale", d, 1000.0, float(vmin), float(vmax)), 0.0, float(lmax / np.float32(2 * 65535.0)))
    e = gpu.get_transform("apply_tone_curve").call_results(nm, gpu.ToneCurve(tone))[0]
    assert isinstance(e, gpu.AflakError) and e.status == gpu.ERR_PANIC


def test_colormap_device_buffer(gpu, orc):
    """The headless stand-in for the GL texture: bytes written to a caller-provided device buffer."""
    img = lognormal((40, 64), 25)
    d = gpu.Image.upload(img)
    out = gpu.Image.alloc((40, 64))  # 40*64*4 bytes = RGBA8
    lut = gpu.Lut("Yellowy")
    gpu.colormap_device(d, lut, 0.2, 4.0, out.device_ptr, channels=4)
    rgba = out.download().view(np.uint8).reshape(40, 64, 4)
    want = orc.make_raw_image(img, 0.2, 4.0, orc.lut_build("Yellowy"), channels=4)
    assert np.array_equal(rgba, want)


# ---- fast display paths (guard-band chain index, exact division by launch-uniform divisors) against the reference's own
# ---- IEEE sequence for every pixel (afk_set_display_exact) on inputs placed at and around every index boundary -------------
def _around(values, ulps):
    """every value together with its neighbours k ulps away, k in `ulps`"""
    v = np.asarray(values, dtype=np.float32)
    b = v.view(np.int32).astype(np.int64)
    out = [(b + k).astype(np.int32).view(np.float32) for k in ulps]
    return np.concatenate(out)


def _pad_image(vals, nx=1024):
    vals = np.asarray(vals, np.float32)
    ny = (vals.size + nx - 1) // nx
    img = np.full(ny * nx, vals[0] if vals.size else 0.0, np.float32)
    img[:vals.size] = vals
    return img.reshape(ny, nx)


SPECIALS = np.array([np.nan, np.inf, -np.inf, 0.0, -0.0, 1e-40, -1e-40, 3e-45, -3.5, 1e30, -1e30, 3.4028235e38, 1.1754944e-38], np.float32)


@pytest.fixture
def exact_switch(gpu):
    yield gpu
    gpu.set_display_exact(False)


@pytest.mark.parametrize("a,tone_len,nm,lims", [
    (1000.0, 256, (0.0, None), (0.2, 0.3, 0.9)),      # BASELINE configs[3]
    (1000.0, 256, (0.0, None), (0.0, 0.5, 1.0)),
    (10.0, 64, (0.0, None), (0.0, 0.5, 1.0)),
    (50000.0, 1024, (0.0, None), (0.1, 0.5, 0.8)),
    (1000.0, 256, (1e-6, 1.3e-5), (0.0, 0.5, 1.0)),    # offset range: indices below zero and beyond the table on both sides
    (300.0, 4096, (0.0, None), (0.0, 0.7, 1.0)),
])
def test_chain_guard_band_matches_exact_sequence_at_every_boundary(exact_switch, orc, a, tone_len, nm, lims):
    """The fast chain takes trunc() of an approximate index and re-evaluates the reference's sequence inside a guard band
    (kernels.cu chain_key_fast).  Inputs sit exactly at, and 1..64 ulps either side of, every tone-index boundary (where a
    wrong guard band would show), plus random and special values; bytes must equal the all-IEEE path's."""
    gpu = exact_switch
    vmin, vmax = np.float32(0.37), np.float32(9.1)
    lna = np.float32(np.log(np.float32(a)))
    lmax = np.float32(np.log(np.float32(a) + np.float32(1.0)) / lna)
    nmin = float(nm[0])
    nmax = float(lmax / np.float32(65535.0)) if nm[1] is None else float(nm[1])
    ts = tone_len - 1
    # invert kf = ((ln(a d + 1) / ln a - nmin) / (nmax - nmin)) ts / 65535 = j for every boundary j (in f64)
    j = np.arange(0, tone_len + 2, dtype=np.float64)
    u = j * 65535.0 / ts * (np.float64(np.float32(nmax) - np.float32(nmin))) + nmin
    t = np.exp(u * np.float64(lna))
    dd = (t - 1.0) / a
    centre = (np.float64(vmin) + dd * np.float64(np.float32(vmax - vmin))).astype(np.float32)
    centre = centre[np.isfinite(centre)]
    rng = np.random.default_rng(int(a) + tone_len)
    vals = np.concatenate([_around(centre, list(range(-64, 65))), SPECIALS,
                           (vmin + (vmax - vmin) * rng.random(200000)).astype(np.float32),
                           np.float32(vmin) - np.exp(rng.standard_normal(2000)).astype(np.float32),
                           np.float32(vmax) + np.exp(rng.standard_normal(2000)).astype(np.float32)])
    img = _pad_image(vals)
    d = gpu.Image.upload(img)
    tone = np.linspace(0.0, ts, tone_len).astype(np.float32) ** np.float32(1.0) if tone_len != 256 else orc.create_curve([(0.0, 0.0), (0.4, 0.7), (1.0, 1.0)])
    tone = np.asarray(tone, np.float32)
    lut = gpu.Lut("Flame")
    lut.set_lims(*lims)
    chain = {"logscale_a": a, "logscale_vmin": float(vmin), "logscale_vmax": float(vmax), "normalize_min": nmin, "normalize_max": nmax,
             "tone_table": tone}
    res = {}
    for exact in (True, False):
        gpu.set_display_exact(exact)
        for ch in (4, 3):
            try:
                res[exact, ch] = gpu.colormap(d, lut, vmin=0.0, vmax=65535.0, channels=ch, chain=chain)
            except gpu.AflakError as e:  # indices beyond the table: the reference panics; both paths must say so
                assert e.status == gpu.ERR_PANIC
                res[exact, ch] = "panic"
    for ch in (4, 3):
        if isinstance(res[True, ch], str) or isinstance(res[False, ch], str):
            assert res[True, ch] == res[False, ch]
            # drop the out-of-range pixels and compare the rest
            keep = img.copy()
            first_oob = _around(centre[min(tone_len, centre.size - 1):][:1], [-200])[0]  # 200 ulps below the boundary of index `len`
            keep[~(keep < min(first_oob, vmax))] = vmin  # also NaN / +Inf
            dk = gpu.Image.upload(keep)
            for exact in (True, False):
                gpu.set_display_exact(exact)
                res[exact, ch] = gpu.colormap(dk, lut, vmin=0.0, vmax=65535.0, channels=ch, chain=chain)
        assert np.array_equal(res[True, ch], res[False, ch]), (ch, int((res[True, ch] != res[False, ch]).any(axis=2).sum()))


@pytest.mark.parametrize("lims", [(0.0, 0.5, 1.0), (0.2, 0.3, 0.9), (0.25, 0.5, 0.75), (0.0, 0.2, 1.0), (0.999, 0.5, 1.0)])
@pytest.mark.parametrize("limits", [(0.37, 9.1), (-3.0, 1e-3), (0.0, 65535.0), (1e-20, 3e-20), (-1e15, 1e15)])
def test_lut_index_fast_division_matches_ieee_at_every_boundary(exact_switch, lims, limits):
    """make_raw_image with the two launch-uniform divisions as div_uniform (kernels.cu lut_index_fast): inputs at and +-2
    ulps around every one of the 65535 index boundaries x = i / 65535."""
    gpu = exact_switch
    vmin, vmax = np.float32(limits[0]), np.float32(limits[1])
    lut = gpu.Lut("HeatMap_Inv")
    lut.set_lims(*lims)
    l0, _, l2 = [np.float64(x) for x in lut.lims()]
    i = np.arange(0, 65536, dtype=np.float64)
    x = i / 65535.0 * (l2 - l0) + l0  # c = i / 65535 (with the default midtone m = c)
    centre = (np.float64(vmin) + x * np.float64(np.float32(vmax - vmin))).astype(np.float32)
    rng = np.random.default_rng(7)
    vals = np.concatenate([_around(centre, [-2, -1, 0, 1, 2]), SPECIALS, (vmin + (vmax - vmin) * rng.random(300000)).astype(np.float32)])
    d = gpu.Image.upload(_pad_image(vals))
    res = {}
    for exact in (True, False):
        gpu.set_display_exact(exact)
        for ch in (4, 3):
            res[exact, ch] = gpu.colormap(d, lut, vmin=float(vmin), vmax=float(vmax), channels=ch)
    for ch in (4, 3):
        assert np.array_equal(res[True, ch], res[False, ch]), (ch, int((res[True, ch] != res[False, ch]).any(axis=2).sum()))


@pytest.mark.parametrize("tone_len", [256, 2, 3, 1000, 4096])
def test_tone_curve_fast_division_matches_ieee_at_every_boundary(exact_switch, orc, tone_len):
    """apply_tone_curve's index is trunc((v * ts) / 65535): the division by the constant as div_uniform (kernels.cu tone_key)."""
    gpu = exact_switch
    from util import assert_bit_equal

    ts = tone_len - 1
    j = np.arange(0, tone_len, dtype=np.float64)
    centre = (j * 65535.0 / ts).astype(np.float32)
    rng = np.random.default_rng(tone_len)
    vals = np.concatenate([_around(centre, list(range(-8, 9))), SPECIALS[SPECIALS < 60000.0], np.array([np.nan], np.float32),
                           (65535.0 * rng.random(300000)).astype(np.float32)])
    vals = vals[~(vals > np.float32(65535.0))]  # larger inputs panic in the reference (covered by test_fused_chain_tone_index_out_of_bounds)
    img = _pad_image(vals)
    d = gpu.Image.upload(img)
    table = (rng.random(tone_len) * ts).astype(np.float32)
    out = {}
    for exact in (True, False):
        gpu.set_display_exact(exact)
        out[exact] = gpu.call("apply_tone_curve", d, gpu.ToneCurve(table)).download()
    assert_bit_equal(out[False], out[True], f"tone curve len {tone_len}")
    assert_bit_equal(out[False], orc.apply_tone_curve(img, table), f"tone curve len {tone_len} vs oracle")
