"""GPU parity at BASELINE.json's full sizes.

C1/C2 (MaNGA 74x74x4563, 100 MB) and C4 (4096x4096 image) are small enough for the oracle: direct parity.
C3 (ALMA 1024x1024x3840, 16.1 GB) is generated on the device by the counter-based generator the oracle shares, so
any sub-block can be regenerated on the host: direct parity on sampled row bands, plus size-independent properties
(additivity over the spectral axis, average over one frame == the frame, fused == unfused, sharded == whole).
"""
import numpy as np
import pytest

from util import assert_bit_equal, assert_reduction, assert_ulp

pytestmark = pytest.mark.gpu

MANGA = (4563, 74, 74)
EW_WINDOWS = ((3134, 3154), (3099, 3119), (3174, 3194))  # src/src/templates.rs:143-463


@pytest.fixture(scope="module")
def manga(gpu, orc):
    nz, ny, nx = MANGA
    host = orc.gen_cube(0xAF1A0001, nz, ny, nx)
    dev = gpu.Image.synthetic(0xAF1A0001, nz, ny, nx)
    return host, dev


def test_device_generator_matches_host_generator(gpu, orc, manga):
    host, dev = manga
    assert_bit_equal(dev.download(), host, "synthetic cube (device vs oracle generator)")


def test_c1_manga_integral_full_range(gpu, orc, manga):
    host, dev = manga
    nz = MANGA[0]
    s64, a64 = orc.sum_f64(host, 0, nz)
    got = gpu.call("integral", dev, 0, nz)[0].download()
    assert_reduction(got, s64, a64, "C1 integral compensated")
    # the compensated result must be at least as close to the f64 sum as the reference's own sequential f32 sum
    ref = orc.integral(host, 0, nz)
    assert np.abs(got - s64).max() <= np.abs(ref - s64).max()
    gpu.set_sum_mode(gpu.SUM_SEQUENTIAL)
    try:
        assert_bit_equal(gpu.call("integral", dev, 0, nz)[0].download(), ref, "C1 integral sequential")
        assert_bit_equal(gpu.call("average", dev, 3134, 3154)[0].download(), orc.average(host, 3134, 3154))
    finally:
        gpu.set_sum_mode(gpu.SUM_COMPENSATED)


def test_c2_manga_equivalent_width_template(gpu, orc, manga):
    host, dev = manga
    on, off1, off2 = [orc.average(host, s, e) for s, e in EW_WINDOWS]
    r2, r1 = orc.ratio_from_bands(3144.0, 3109.0, 3184.0)
    want = orc.equivalent_width(orc.linear_composition(off1, off2, r2, r1), on, 20.0, 1e10, True)
    gpu.set_sum_mode(gpu.SUM_SEQUENTIAL)
    try:
        img, (mn, mx) = gpu.fused_equivalent_width(dev, *EW_WINDOWS, 20.0, 1e10, True, want_minmax=True)
        assert_bit_equal(img.download(), want, "C2 fused EW (sequential)")
        assert (mn, mx) == orc.image_min_max(want)
    finally:
        gpu.set_sum_mode(gpu.SUM_COMPENSATED)
    comp = gpu.fused_equivalent_width(dev, *EW_WINDOWS, 20.0, 1e10, True).download()
    np.testing.assert_allclose(comp, want, rtol=2e-4, atol=2e-5)  # both are f32 evaluations of an ill-conditioned quotient
    # linear_composition of three slices (two chained calls)
    f = [gpu.call("slice_one_frame", dev, k) for k in (3109, 3144, 3184)]
    out = gpu.call("linear_composition", gpu.call("linear_composition", f[0], f[1], 0.5, 0.25), f[2], 1.0, 0.25).download()
    assert_bit_equal(out, orc.linear_composition(orc.linear_composition(host[3109], host[3144], 0.5, 0.25), host[3184], 1.0, 0.25))


def test_c4_image_chain_4096(gpu, orc):
    rng = np.random.default_rng(41)
    img = np.exp(rng.standard_normal((4096, 4096))).astype(np.float32)
    img[rng.random(img.shape) < 1e-3] = np.nan
    d = gpu.Image.upload(img)
    mn, mx = gpu.call("image_min_max", d)
    assert (mn, mx) == orc.image_min_max(img)
    assert d.vmin_vmax() == (orc.get_vmin(img), orc.get_vmax(img))
    for gradient, lims in (("Flame", (0.0, 0.5, 1.0)), ("Flame", (0.2, 0.3, 0.9))):
        lut = gpu.Lut(gradient)
        lut.set_lims(*lims)
        table = orc.lut_build(gradient)
        for ch in (4, 3):
            got = gpu.colormap(d, lut, channels=ch)
            want = orc.make_raw_image(img, orc.get_vmin(img), orc.get_vmax(img), table, lims=lut.lims(), channels=ch, threads=orc.max_threads())
            assert np.array_equal(got, want), int((got != want).sum())
    # down_sampling at full size
    assert_bit_equal(gpu.call("down_sampling", d, 4).download(), img[::4, ::4])
    # tone-curve index at full size: bit-exact
    scaled = gpu.call("normalize_image", d, 0.0, float(mx) / 65535.0)
    host_scaled = (img / (np.float32(float(mx) / 65535.0) - np.float32(0.0))).astype(np.float32)
    assert_bit_equal(scaled.download(), host_scaled)
    table = orc.create_curve([(0.0, 0.0), (0.4, 0.7), (1.0, 1.0)])
    ok = np.nan_to_num(host_scaled, nan=0.0) <= 65535.0
    assert ok.all()
    assert_bit_equal(gpu.call("apply_tone_curve", scaled, gpu.ToneCurve(table)).download(), orc.apply_tone_curve(host_scaled, table))


ALMA = (3840, 1024, 1024)


@pytest.fixture(scope="module")
def alma(gpu):
    nz, ny, nx = ALMA
    return gpu.Image.synthetic(0xAF1A0003, nz, ny, nx)


def test_c3_alma_sampled_bands_against_oracle(gpu, orc, alma):
    nz, ny, nx = ALMA
    half = nz // 2
    fused = gpu.fused_average_ratio(alma, 0, half, half, nz).download()
    a0 = gpu.call("average", alma, 0, half)[0]
    a1 = gpu.call("average", alma, half, nz)[0]
    unfused = gpu.call("ratio", a0, a1).download()
    # compensated mode: the fused launch (two windows) splits the spectral axis between CTAs at other frames than the
    # two single-window launches, so the f64 combination of the Kahan partials can round the other way: <= 1 ulp on each
    # mean (values in [2,4)), which is up to 2 x 1.5 ulp of a quotient just below 1, plus its own rounding: <= 4 ulp.  (Sequential mode is bit-identical: test_c3_alma_properties and tests/test_gpu_graph.py.)
    assert_ulp(fused, unfused, 4, "C3 fused vs operator-by-operator")
    h0, h1 = a0.download(), a1.download()
    for y0 in (0, 513, 1020):
        band = orc.gen_cube(0xAF1A0003, nz, ny, nx, y0=y0, rows=4)  # regenerate 4 rows of the 16 GB cube on the host
        s0, ab0 = orc.sum_f64(band, 0, half)
        s1, ab1 = orc.sum_f64(band, half, nz)
        assert_reduction(h0[y0:y0 + 4], s0, ab0, f"C3 average[0,{half}) rows {y0}..", divisor=float(half))
        assert_reduction(h1[y0:y0 + 4], s1, ab1, f"C3 average[{half},{nz}) rows {y0}..", divisor=float(nz - half))
        np.testing.assert_allclose(fused[y0:y0 + 4], s0 / s1, rtol=1e-5)


def test_c3_alma_properties(gpu, orc, alma):
    nz, ny, nx = ALMA
    whole = gpu.call("integral", alma, 0, nz)[0].download().astype(np.float64)
    a, b = 1234, 2901
    parts = sum(gpu.call("integral", alma, s, e)[0].download().astype(np.float64) for s, e in ((0, a), (a, b), (b, nz)))
    # additivity over the spectral axis: each piece is within 1e-5 of its exact sum (positive data), so is the total
    np.testing.assert_allclose(parts, whole, rtol=3e-5)
    # average over one frame is the frame, bit for bit; so is slice_one_frame (zero-copy view)
    frame = gpu.call("slice_one_frame", alma, 2000).download()
    assert_bit_equal(gpu.call("average", alma, 2000, 2001)[0].download(), frame)
    band = orc.gen_cube(0xAF1A0003, nz, ny, nx, y0=700, rows=2)
    assert_bit_equal(frame[700:702], band[2000])
    # sequential mode at full size on sampled rows: bit-exact with the reference's left-to-right f32 sum
    gpu.set_sum_mode(gpu.SUM_SEQUENTIAL)
    try:
        seq = gpu.call("integral", alma, 0, nz)[0].download()
    finally:
        gpu.set_sum_mode(gpu.SUM_COMPENSATED)
    assert_bit_equal(seq[700:702], orc.integral(band, 0, nz), "C3 sequential integral rows 700..702")


def test_c5_sharded_bands_equal_whole(gpu, orc, alma):
    """Sharding is exact: reducing the y-bands separately (as each rank does) and laying them side by side gives the
    same image as reducing the whole cube.  Emulated on one GPU with band cubes from the shared generator.
    Sequential mode: same bits.  Compensated mode: the split of the spectral axis between CTAs depends on the band
    size, so the f64 combination order differs -> equal to within one f32 ulp."""
    nz, ny, nx = 512, 4096, 512
    world = 8
    whole_dev = gpu.Image.synthetic(0xAF1A0005, nz, ny, nx)
    whole = {}
    for mode in (gpu.SUM_SEQUENTIAL, gpu.SUM_COMPENSATED):
        gpu.set_sum_mode(mode)
        whole[mode] = gpu.call("integral", whole_dev, 0, nz)[0].download()
    del whole_dev
    try:
        for rank in (0, 3, 7):
            y0, rows = gpu.shard_band(ny, world, rank)
            band = gpu.Image.synthetic(0xAF1A0005, nz, ny, nx, y0=y0, rows=rows)
            gpu.set_sum_mode(gpu.SUM_SEQUENTIAL)
            assert_bit_equal(gpu.call("integral", band, 0, nz)[0].download(), whole[gpu.SUM_SEQUENTIAL][y0:y0 + rows], f"band of rank {rank}")
            gpu.set_sum_mode(gpu.SUM_COMPENSATED)
            got = gpu.call("integral", band, 0, nz)[0].download()
            np.testing.assert_allclose(got, whole[gpu.SUM_COMPENSATED][y0:y0 + rows], rtol=1.2e-7)
    finally:
        gpu.set_sum_mode(gpu.SUM_COMPENSATED)


# ---- SURVEY 8f rows at full size -----------------------------------------------------------------------------------

def test_c3_alma_axis0_operators_sampled_bands(gpu, orc, alma):
    """variance / stddev / arg maps / fold / centroid / median over the whole 16 GB cube, checked against the oracle on
    row bands regenerated on the host, plus properties that hold at any size."""
    nz, ny, nx = ALMA
    alma.set_axis(2, "WAVE", "Angstrom")
    wcs = (3621.6, 1.0, 0.83)
    alma.set_wcs_axis(2, *wcs)
    var = gpu.call("variance", alma, 0, nz).download()
    std = gpu.call("stddev", alma, 0, nz).download()
    arg, peak = (x.download() for x in gpu.call("extract_argmin_max_wavelength", alma, 0, nz, False))
    lo, hi = (x.download() for x in gpu.call("peak_based_wavelength_range", alma, 0, nz, 5, False))
    cen = gpu.call("extract_centrobaric_wavelength", alma, 0, nz).download()
    med = gpu.call("median", alma, 100, 1125).download()   # odd window of 1025 frames
    mn = gpu.call("extract_argmin_max_wavelength", alma, 0, nz, True)[1].download()
    # properties: stddev^2 == variance; min <= median <= max; the two index maps are 10 apart; the peak wavelength is the
    # world coordinate of an integer frame inside the range
    np.testing.assert_allclose(std.astype(np.float64) ** 2, var, rtol=1e-6)
    assert np.all(mn <= med) and np.all(med <= peak)
    assert np.array_equal(hi - lo, np.full_like(hi, 10.0))
    k = (arg.astype(np.float64) - wcs[0]) / wcs[2]
    assert np.all(np.abs(k - np.round(k)) < 1e-2) and k.min() >= -0.01 and k.max() <= nz - 1 + 0.01
    assert np.array_equal(np.round(k), hi - 5.0)
    for y0 in (0, 513, 1021):
        band = orc.gen_cube(0xAF1A0003, nz, ny, nx, y0=y0, rows=3)
        rows = slice(y0, y0 + 3)
        np.testing.assert_allclose(var[rows], orc.variance_f64(band, 0, nz), rtol=1e-5)
        assert_bit_equal(arg[rows], orc.argmap(band, 0, nz, 0, False, wcs=wcs), f"C3 argmax rows {y0}")
        assert_bit_equal(peak[rows], orc.minmax_fold(band, 0, nz, False), f"C3 max fold rows {y0}")
        assert_bit_equal(lo[rows], orc.argmap(band, 0, nz, -5, False), f"C3 peak range rows {y0}")
        np.testing.assert_allclose(cen[rows], orc.centroid_f64(band, 0, nz, wcs), rtol=1e-5)
        assert_bit_equal(med[rows], orc.median(band, 100, 1125), f"C3 median rows {y0}")
    # sequential mode on the same rows: the reference's own bits (Welford f32 with FMA)
    gpu.set_sum_mode(gpu.SUM_SEQUENTIAL)
    try:
        seq = gpu.call("variance", alma, 0, nz).download()
    finally:
        gpu.set_sum_mode(gpu.SUM_COMPENSATED)
    assert_bit_equal(seq[513:516], orc.variance(orc.gen_cube(0xAF1A0003, nz, ny, nx, y0=513, rows=3), 0, nz), "C3 Welford rows 513..516")


def test_c3_alma_extract_wave_checksum_of_checksums(gpu, orc, alma):
    """Two orthogonal reductions of the same 4 G voxels must agree: the sum over frames of extract_wave(All) equals the
    sum over pixels of integral (both within 1e-5 of the exact total, positive data); spot frames against numpy."""
    nz, ny, nx = ALMA
    wave = gpu.call("extract_wave", alma, gpu.Roi()).download().astype(np.float64)
    image = gpu.call("integral", alma, 0, nz)[0].download().astype(np.float64)
    assert wave.shape == (nz,)
    np.testing.assert_allclose(wave.sum(), image.sum(), rtol=2e-5)
    for k in (0, 1999, nz - 1):
        frame = gpu.call("slice_one_frame", alma, k).download().astype(np.float64)
        np.testing.assert_allclose(wave[k], frame.sum(), rtol=1e-5)
    # sequential mode: every frame summed in the reference's own order (roi.rs:17-34); sampled frames bit for bit
    gpu.set_sum_mode(gpu.SUM_SEQUENTIAL)
    try:
        seq = gpu.call("extract_wave", alma, gpu.Roi()).download()
    finally:
        gpu.set_sum_mode(gpu.SUM_COMPENSATED)
    for k in (0, 31, 32, 1999, nz - 1):
        frame = gpu.call("slice_one_frame", alma, k).download()
        assert_bit_equal(seq[k:k + 1], orc.extract_wave(frame[None]), f"C3 sequential extract_wave frame {k}")
    np.testing.assert_allclose(seq, wave, rtol=5e-3)  # a million f32 additions in sequence: ~6e-4 off the f64 sum
    # a box region: the waveform template at full size, against the host generator
    box = [(x, y) for y in range(700, 703) for x in range(100, 164)]
    got = gpu.call("extract_wave", alma, gpu.Roi(box)).download()
    band = orc.gen_cube(0xAF1A0003, nz, ny, nx, y0=700, rows=3)
    want = band[:, :, 100:164].astype(np.float64).sum(axis=(1, 2))
    np.testing.assert_allclose(got, want, rtol=1e-5)


def test_c3_size_ingest_round_trip(gpu, orc):
    """A 2 GiB BITPIX 16 data unit (1 Gi voxels) through the chunked pipeline: 32 chunks, both streams, staging ring.
    Property: decode(encode(x)) == x * bscale + bzero for every voxel, checked on the device; samples against the oracle."""
    n = 1 << 30
    rng = np.random.default_rng(41)
    block = rng.integers(-32768, 32768, 1 << 24).astype(">i2")
    raw = np.tile(block, n // block.size)
    img = gpu.ingest(raw, gpu.FITS_I16, (1024, 1024, 1024), bscale=0.5, bzero=100.0)
    want_block = orc.fits_decode(block.tobytes(), 2, block.size, bscale=0.5, bzero=100.0)
    got = img.download().reshape(-1)
    for off in (0, 17 * block.size, n - block.size):
        assert np.array_equal(got[off:off + block.size], want_block)
    # extremes of the whole decoded cube follow from the encoding
    mn, mx = gpu.call("image_min_max", img)
    assert float(mn) == float(want_block.min()) and float(mx) == float(want_block.max())


def test_c4_size_display_statistics(gpu, orc):
    """4096 x 4096: histogram counts sum to the pixel count and match the oracle; vmed / vmad idempotent under re-upload."""
    rng = np.random.default_rng(42)
    a = rng.lognormal(0.0, 1.0, (4096, 4096)).astype(np.float32)
    a[::997, ::13] = np.nan
    img = gpu.Image.upload(a)
    vmin, vmax = img.vmin_vmax()
    counts, _, _ = img.histogram(float(vmin), float(vmax))
    assert counts.sum() == a.size and np.array_equal(counts, orc.histogram(a, float(vmin), float(vmax))[0])
    z = img.zscale()
    zo = orc.zscale(a)
    assert z[0].tobytes() == zo[0].tobytes() and z[1].tobytes() == zo[1].tobytes()
