"""GPU parity: display statistics -- histogram, normalised median / MAD, zscale.

Reference: aflak_plot/src/imshow/hist.rs:10-66, lims.rs:37-106, zscale.rs:45-242, image.rs:155-176.
Tier A: counts, order statistics and zscale limits are bit-identical to the oracle's restatement.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def images(orc):
    rng = np.random.default_rng(21)
    yield "lognormal 257x300", rng.lognormal(0.0, 1.0, (257, 300)).astype(np.float32)
    a = rng.normal(100.0, 15.0, (512, 512)).astype(np.float32)
    a[::37, ::11] = np.nan
    yield "normal 512x512 with NaN", a.copy()
    a[5, 5] = np.inf
    yield "normal 512x512 with NaN and +Inf", a.copy()
    a[6, 6] = -np.inf
    yield "normal 512x512 with NaN and +-Inf", a
    yield "cube frame 74x74", orc.gen_cube(0xAF1A0021, 8, 74, 74)[3]
    b = np.round(rng.normal(0, 3, (129, 67))).astype(np.float32)  # heavy ties
    yield "quantised 129x67", b
    yield "single row 1x1000", rng.uniform(-1, 1, (1, 1000)).astype(np.float32)
    yield "tiny 2x2", np.array([[1.0, 2.0], [4.0, 8.0]], np.float32)
    yield "even count 4x5", rng.uniform(0, 1, (4, 5)).astype(np.float32)


def test_histogram_counts_exact(gpu, orc):
    for name, a in images(orc):
        img = gpu.Image.upload(a)
        fin = a[np.isfinite(a)]
        for vmin, vmax in [(float(fin.min()), float(fin.max())), (float(np.median(fin)), float(fin.max())), (0.0, 1.0)]:
            got, s0, e0 = img.histogram(vmin, vmax)
            want, s1, e1 = orc.histogram(a, vmin, vmax)
            assert np.array_equal(got, want), f"{name} [{vmin},{vmax}]"
            assert got.sum() == a.size
            assert np.array_equal(s0, s1) and np.array_equal(e0, e1)


def test_histogram_color(gpu, orc):
    rng = np.random.default_rng(22)
    rgb = rng.uniform(-100, 70000, (3, 211, 130)).astype(np.float32)
    rgb[1, ::7] = np.nan
    got, s0, e0 = gpu.Image.upload(rgb).histogram(0.0, 65535.0, color=True)
    want, s1, e1 = orc.histogram(rgb, 0.0, 65535.0, color=True)
    assert got.shape == (256, 3) and np.array_equal(got, want) and np.array_equal(s0, s1) and np.array_equal(e0, e1)
    e = None
    try:
        gpu.Image.upload(rgb[:2]).histogram(0.0, 1.0, color=True)
    except gpu.AflakError as err:
        e = err
    assert e is not None and e.status == gpu.ERR_SHAPE


def test_vmed_vmad_bit_exact(gpu, orc):
    for name, a in images(orc):
        try:
            want = orc.vmed_vmad(a)
        except orc.OracleError as err:  # vmin = -inf / vmax = +inf: every normalised value is NaN -> "Empty data!, med"
            with pytest.raises(gpu.AflakError) as e:
                gpu.Image.upload(a).vmed_vmad()
            assert err.msg in str(e.value), name
            continue
        got = gpu.Image.upload(a).vmed_vmad()
        assert got[0].tobytes() == want[0].tobytes() and got[1].tobytes() == want[1].tobytes(), f"{name}: {got} != {want}"


def test_vmed_vmad_full_size_c4(gpu, orc):
    """4096 x 4096 (BASELINE config 4 size): the select runs over 16.7 M pixels; checked against numpy's partition."""
    rng = np.random.default_rng(23)
    a = rng.lognormal(0.0, 1.0, (4096, 4096)).astype(np.float32)
    a[::1000, ::3] = np.nan
    got = gpu.Image.upload(a).vmed_vmad()
    vmin, vmax = np.nanmin(a), np.nanmax(a)
    d = ((a - vmin) / (vmax - vmin)).reshape(-1)
    d = d[np.isfinite(d)]
    mid = d.size // 2
    part = np.partition(d, [mid - 1, mid])
    vmed = (part[mid] + part[mid - 1]) / np.float32(2.0) if d.size % 2 == 0 else part[mid]
    e = np.abs(d - vmed)
    part = np.partition(e, [mid - 1, mid])
    vmad = (part[mid] + part[mid - 1]) / np.float32(2.0) if d.size % 2 == 0 else part[mid]
    assert got[0] == np.float32(vmed) and got[1] == np.float32(vmad)


def big_images():
    """Megapixel images: vmed / vmad take the one-pass (sample-bracketed) select; the patterns stress its bracket, its bins and
    its fall-back to the radix select."""
    rng = np.random.default_rng(24)
    a = rng.normal(100.0, 15.0, (1024, 1024)).astype(np.float32)
    yield "normal 1024x1024, even count", a.copy()
    a[::37, ::11] = np.nan
    a[5, 5] = 0.0
    yield "normal with NaN (odd or even kept count)", a.copy()
    a[7, 7] = np.nan
    yield "normal with one more NaN (the other parity)", a.copy()
    yield "quantised 1200x1000 (heavy ties: one key value in the bracket)", np.round(rng.normal(0, 3, (1200, 1000))).astype(np.float32)
    yield "quantised, wide (several tied values in the bracket)", np.round(rng.normal(0, 300, (1100, 1000))).astype(np.float32)
    b = rng.lognormal(0.0, 1.0, (1500, 1100)).astype(np.float32)
    b[rng.random(b.shape) < 0.3] = 0.0
    yield "30 % zeros + lognormal", b
    yield "two values", np.where(rng.random((1024, 1100)) < 0.5, np.float32(1.0), np.float32(7.0)).astype(np.float32)
    yield "ramp (sorted image)", np.arange(1024 * 1024, dtype=np.float32).reshape(1024, 1024) * np.float32(0.37)
    c = rng.uniform(-1.0, 1.0, (1024, 1024)).astype(np.float32)
    c[0, 0] = -1e30
    c[0, 1] = 1e30
    yield "outliers stretch the range (all normalised values within a few ulps)", c
    d = rng.normal(0.0, 1.0, (1025, 1023)).astype(np.float32)
    d[d > 0.0] = np.float32(0.5)  # the upper half is one value: the median sits at the edge of a huge tie
    yield "upper half tied", d
    e = rng.normal(10.0, 2.0, (1024, 1024)).astype(np.float32)
    e[3, 3] = np.inf  # vmax = +Inf: no raw-key shortcut, every finite pixel normalises to 0, the Inf pixel to NaN (dropped)
    yield "one +Inf pixel (infinite range: one key value in the bracket, normalised path for both statistics)", e
    f = rng.normal(0.0, 1.0, (1024, 1024)).astype(np.float32)
    f[::2, :] = np.nan  # half of every warp's stream is dropped
    yield "every other row NaN", f


def test_vmed_vmad_one_pass_bit_exact(gpu, orc):
    for name, a in big_images():
        want = orc.vmed_vmad(a)
        got = gpu.Image.upload(a).vmed_vmad()
        assert got[0].tobytes() == want[0].tobytes() and got[1].tobytes() == want[1].tobytes(), f"{name}: {got} != {want}"
    for a in [np.full((1024, 1024), 2.0, np.float32), np.full((1024, 1030), np.nan, np.float32)]:
        with pytest.raises(gpu.AflakError) as e:
            gpu.Image.upload(a).vmed_vmad()
        assert e.value.status == gpu.ERR_UNEXPECTED_INPUT and "Empty data!, med" in str(e.value)


def test_vmed_vmad_errors(gpu, orc):
    for a in [np.full((5, 5), 2.0, np.float32), np.full((3, 3), np.nan, np.float32)]:
        with pytest.raises(gpu.AflakError) as e:
            gpu.Image.upload(a).vmed_vmad()
        assert e.value.status == gpu.ERR_UNEXPECTED_INPUT and "Empty data!, med" in str(e.value)


def test_zscale_bit_exact(gpu, orc):
    rng = np.random.default_rng(24)
    cases = list(images(orc))
    big = rng.lognormal(0.0, 1.0, (1500, 2100)).astype(np.float32)
    big[::50] = np.nan
    cases.append(("lognormal 1500x2100, NaN rows", big))
    star = rng.normal(10.0, 1.0, (600, 600)).astype(np.float32)
    star[280:320, 280:320] += 500.0  # a bright source: the iterative rejection kicks in
    cases.append(("gaussian + source", star))
    for name, a in cases:
        img = gpu.Image.upload(a)
        for nsamples, contrast in [(1000, 0.25), (1000, 1.0), (100, 0.25), (37, -1.0), (5, 0.25)]:
            got = img.zscale(nsamples, contrast)
            want = orc.zscale(a, nsamples, contrast)
            if want[0] is None:
                assert got == (None, None), name
            else:
                assert got[0].tobytes() == want[0].tobytes() and got[1].tobytes() == want[1].tobytes(), f"{name} ({nsamples},{contrast}): {got} != {want}"
    assert gpu.Image.upload(np.full((6, 6), np.nan, np.float32)).zscale() == (None, None)
    assert gpu.Image.upload(cases[0][1]).zscale(0, 0.25) == (None, None)
    with pytest.raises(gpu.AflakError):  # zsc_sample fails on anything but 2-D and the reference unwraps it
        gpu.Image.upload(np.zeros((2, 3, 4), np.float32)).zscale()
