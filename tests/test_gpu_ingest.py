"""GPU parity: FITS data unit -> device-resident f32 image.

Reference: aflak_primitives/src/unit.rs:283-331 (WcsArray::from_hdu), fits.rs:42-103 (FitsDataToArray), lib.rs:826-910
(run_open_fits, run_fits_to_image); fitrs 0.5.0 read_data restated from the FITS standard.
Tier A: the decoded image is bit-identical to the oracle's byte-by-byte restatement (NaN positions exact).
"""
import json
import os

import numpy as np
import pytest

from util import assert_bit_equal, fits_hdu_bytes

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
FORMATS = [(0, ">f4"), (1, ">f8"), (2, ">i2"), (3, ">i4"), (4, ">u4")]


def raw_array(rng, dt, shape):
    n = int(np.prod(shape))
    if dt == ">f4":
        a = rng.normal(0, 100, n).astype(np.float32)
        a[::97] = np.nan
        a[1::101] = np.inf
        a[2::103] = -0.0
        a[3::107] = 1e-42  # denormal
    elif dt == ">f8":
        a = rng.normal(0, 1e3, n)
        a[::97] = np.nan
        a[1::101] = 1e300   # overflows f32 -> inf
        a[2::103] = 1e-300  # underflows -> 0
        a[3::107] = 16777217.0  # ties-to-even at the f32 grid
    elif dt == ">i2":
        a = rng.integers(-32768, 32768, n)
    elif dt == ">i4":
        a = rng.integers(-2 ** 31, 2 ** 31, n)
        a[::5] = rng.integers(-1000, 1000, len(a[::5]))
    else:
        a = rng.integers(0, 2 ** 32, n)
    return a.astype(dt).reshape(shape)


@pytest.mark.parametrize("fmt,dt", FORMATS)
@pytest.mark.parametrize("shape", [(7,), (33, 5), (4, 9, 11), (2, 3, 8, 8), (1000, 257)])
def test_ingest_memory_all_formats(gpu, orc, fmt, dt, shape):
    rng = np.random.default_rng(fmt * 10 + len(shape))
    a = raw_array(rng, dt, shape)
    n = a.size
    for blank, bscale, bzero in [(None, 1.0, 0.0), (int(a.reshape(-1)[3]) if dt[1] != "f" else None, 0.25, 32768.0), (None, 1e-3, -7.5)]:
        img = gpu.ingest(a.tobytes(), fmt, shape, blank=blank, bscale=bscale, bzero=bzero)
        assert img.shape == tuple(shape)
        want = orc.fits_decode(a.tobytes(), fmt, n, blank=blank, bscale=bscale, bzero=bzero).reshape(shape)
        got = img.download()
        assert_bit_equal(got, want, f"ingest {dt}{shape} blank={blank}")
        assert np.array_equal(np.signbit(got[got == 0]), np.signbit(want[want == 0]))
        if blank is not None:
            assert np.isnan(got.reshape(-1)[3])


@pytest.mark.parametrize("fmt,dt", FORMATS)
def test_ingest_multi_chunk_pipeline(gpu, orc, fmt, dt):
    """More than two 64 MiB chunks: both pipeline streams, both staging buffers and the ragged tail are exercised."""
    es = int(dt[2])
    n = (3 * (64 << 20) + 4096 * 13 + 20) // es  # not a multiple of 4 elements per chunk boundary
    rng = np.random.default_rng(7)
    a = raw_array(rng, dt, (n,))
    img = gpu.ingest(a, fmt, (n,), bscale=0.5, bzero=1.0)
    want = orc.fits_decode(a.view(np.uint8), fmt, n, bscale=0.5, bzero=1.0)
    assert_bit_equal(img.download(), want, f"ingest pipeline {dt}")
    st = gpu.ingest_stats()
    assert st["bytes"] == n * es and st["device_ms"] > 0


@pytest.mark.parametrize("fmt,dt", [(0, ">f4"), (2, ">i2"), (1, ">f8")])
def test_ingest_band_matches_full(gpu, orc, fmt, dt):
    nz, ny, nx = 37, 50, 24
    a = raw_array(np.random.default_rng(9), dt, (nz, ny, nx))
    full = orc.fits_decode(a.tobytes(), fmt, a.size, bscale=2.0, bzero=-1.0).reshape(nz, ny, nx)
    for world in (2, 3, 8):
        rows_seen = 0
        for rank in range(world):
            y0, rows = gpu.shard_band(ny, world, rank)
            img = gpu.ingest(a.tobytes(), fmt, (nz, ny, nx), bscale=2.0, bzero=-1.0, band=(y0, rows))
            assert img.shape == (nz, rows, nx)
            assert_bit_equal(img.download(), full[:, y0:y0 + rows], f"band {rank}/{world} {dt}")
            rows_seen += rows
        assert rows_seen == ny


def test_fits_to_image_reference_crop(gpu, orc):
    """The reference's own data fixture (cropped): decoded image == the committed numpy decode, metadata from the header."""
    fits = gpu.call("open_fits", gpu.Paths([os.path.join(GOLD, "test_fits_crop.fits")]), 0)
    img = gpu.call("fits_to_image", fits, 0, "")
    crop = np.load(os.path.join(GOLD, "test_fits_crop.npy"))
    facts = json.load(open(os.path.join(GOLD, "test_fits_facts.json")))
    assert img.shape == crop.shape == (200, 12, 12)
    assert_bit_equal(img.download(), crop * np.float32(1.0) + np.float32(0.0), "test.fits crop")
    assert img.unit == facts["bunit"] and img.has_meta and img.tag is None
    assert [img.axis(i)[0] for i in range(3)] == facts["ctype"] and img.axis(3) == ("", "")
    # linear WCS of the spectral axis: CRVAL3 -199, CRPIX3 1, CDELT3 2 -> frame k is at -199 + 2k km/s
    assert img.pix2world(2, 0.0) == -199.0 and img.pix2world(2, 100.0) == 1.0
    # the image is a normal cube for the rest of the path
    s64 = np.array(facts["crop_sum_f64_40_160"])
    got = gpu.call("integral", img, 40, 160)[0].download()
    assert np.allclose(got, s64, rtol=1e-5, atol=0)
    # chained through the graph evaluator like the reference's default template (open_fits -> fits_to_image -> integral)
    g = gpu.Graph()
    c_path = g.add_constant(gpu.Paths([os.path.join(GOLD, "test_fits_crop.fits")]))
    t_open, t_img, t_int = g.add_transform("open_fits"), g.add_transform("fits_to_image"), g.add_transform("integral")
    c_a, c_b = g.add_constant(40), g.add_constant(160)
    g.connect(c_path, 0, t_open, 0)
    g.connect(t_open, 0, t_img, 0)
    g.connect(t_img, 0, t_int, 0)
    g.connect(c_a, 0, t_int, 1)
    g.connect(c_b, 0, t_int, 2)
    assert np.array_equal(g.compute(t_int, 0).download(), got)


def test_fits_to_image_extensions_scaling_and_bands(gpu, orc, tmp_path):
    rng = np.random.default_rng(11)
    sci = raw_array(rng, ">i2", (40, 30, 16))
    sci.reshape(-1)[5] = -7
    err = raw_array(rng, ">f8", (30, 16))
    cnt = raw_array(rng, ">i4", (9, 30, 16))
    blob = (fits_hdu_bytes(None, [("OBJECT", "synthetic")]) +
            fits_hdu_bytes(sci, [("BSCALE", 0.5), ("BZERO", 32768), ("BLANK", -7), ("BUNIT", "adu"), ("CTYPE3", "WAVE"), ("CUNIT3", "Angstrom"),
                                 ("CRVAL3", 3621.6), ("CRPIX3", 1), ("CDELT3", 0.83), ("CRVAL2", 10.0), ("CRPIX2", 15.5), ("CDELT2", 0.1)],
                           primary=False, extname="SCI") +
            fits_hdu_bytes(err, [], primary=False, extname="ERR") +
            fits_hdu_bytes(cnt, [("BLANK", int(cnt.reshape(-1)[2]))], primary=False))
    p = tmp_path / "multi.fits"
    p.write_bytes(blob)
    fits = gpu.Fits.open(p)
    want_sci = orc.fits_decode(sci.tobytes(), 2, sci.size, blank=-7, bscale=0.5, bzero=32768.0).reshape(sci.shape)
    for args in [(0, "SCI"), (1, ""), (1, "NOPE")]:  # by name first, else by index (lib.rs:889-891)
        img = gpu.call("fits_to_image", fits, *args)
        assert_bit_equal(img.download(), want_sci, f"SCI via {args}")
    assert np.isnan(want_sci.reshape(-1)[5]) and img.unit == "adu" and img.axis(2) == ("WAVE", "Angstrom")
    assert img.pix2world(2, 10.0) == orc.pix2world((3621.6, 1.0, 0.83), 10.0)
    assert_bit_equal(gpu.call("fits_to_image", fits, 0, "ERR").download(), orc.fits_decode(err.tobytes(), 1, err.size).reshape(err.shape), "ERR")
    img = gpu.call("fits_to_image", fits, 3, "")
    assert_bit_equal(img.download(), orc.fits_decode(cnt.tobytes(), 3, cnt.size, blank=int(cnt.reshape(-1)[2])).reshape(cnt.shape), "HDU 3")
    assert img.unit == "" and img.axis(0) == ("", "")
    # primary HDU without data
    e = gpu.get_transform("fits_to_image").call_results(fits, 0, "")[0]
    assert isinstance(e, gpu.AflakError) and e.status == gpu.ERR_FITS
    # per-rank band reads: every GPU reads only its own rows, world coordinates are unchanged
    for world in (2, 4):
        for rank in range(world):
            y0, rows = gpu.shard_band(30, world, rank)
            band = fits.to_image(0, "SCI", world=world, rank=rank)
            assert_bit_equal(band.download(), want_sci[:, y0:y0 + rows], f"file band {rank}/{world}")
            assert band.pix2world(1, 0.0) == img_world(orc, y0)
    # an 8-bit HDU is FitsData::Characters: unsupported (unit.rs:291-293)
    p2 = tmp_path / "chars.fits"
    p2.write_bytes(fits_hdu_bytes(np.zeros((4, 4), dtype="u1"), []))
    e = gpu.get_transform("fits_to_image").call_results(gpu.Fits.open(p2), 0, "")[0]
    assert e.status == gpu.ERR_FITS and e.message == "Unsupported data array format: 'Characters'."


def img_world(orc, y0):
    return orc.pix2world((10.0, 15.5, 0.1), float(y0))
