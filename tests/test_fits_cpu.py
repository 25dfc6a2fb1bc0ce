"""CPU tests of the FITS ingest front end: header parsing, HDU lookup and the reference's error wording work
without a GPU; the decode itself needs one (tests/test_gpu_ingest.py) and fails loudly here.  Also pins the oracle's
restatement of the decode arithmetic with hand-derived values.

Reference: aflak_primitives/src/lib.rs:145-170, 826-910; unit.rs:244-331; fits.rs:42-103.
"""
import json
import os

import numpy as np
import pytest

from util import fits_hdu_bytes

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_open_reference_crop_headers(afk):
    f = afk.Fits.open(os.path.join(GOLD, "test_fits_crop.fits"))
    facts = json.load(open(os.path.join(GOLD, "test_fits_facts.json")))
    assert len(f) == 1
    assert f.header("BITPIX") == str(facts["bitpix"]) and f.header("NAXIS") == "3"
    assert [f.header(f"NAXIS{i}") for i in (1, 2, 3)] == ["12", "12", "200"]
    assert f.header("BUNIT") == facts["bunit"]                     # 'K                 ' -> right-trimmed
    assert [f.header(f"CTYPE{i}") for i in (1, 2, 3)] == facts["ctype"]
    assert float(f.header("BSCALE")) == facts["bscale"] and float(f.header("BZERO")) == facts["bzero"]
    assert f.header("NOSUCHKEY") is None and f.header("END") is None


def test_open_fits_transform_and_errors(afk, tmp_path):
    t = afk.get_transform("open_fits")
    assert t.version == (1, 0, 0) and t.inputs == [("path", "Paths", None), ("n", "Integer", 0)] and t.outputs == ["Fits"]
    t2 = afk.get_transform("fits_to_image")
    assert t2.inputs == [("fits", "Fits", None), ("hdu_idx", "Integer", 0), ("extension", "Str", "")] and t2.outputs == ["Image"]
    crop = os.path.join(GOLD, "test_fits_crop.fits")
    fits = afk.call("open_fits", afk.Paths([crop]), 0)
    assert isinstance(fits, afk.Fits) and len(fits) == 1
    # precheck!(pathlist_len > n) (lib.rs:829)
    e = t.call_results(afk.Paths([crop]), 1)[0]
    assert isinstance(e, afk.AflakError) and e.status == afk.ERR_UNEXPECTED_INPUT and e.message == "Expected pathlist_len > n, but got 1 > 1"
    e = t.call_results(afk.Paths([]), 0)[0]
    assert e.message == "Expected pathlist_len > n, but got 0 > 0"
    # IOErr::IoError(err, "Could not open file {:?}")
    missing = str(tmp_path / "nope.fits")
    e = t.call_results(afk.Paths([missing]), 0)[0]
    assert e.status == afk.ERR_IO and e.message == f'Could not open file "{missing}". This was caused by \'No such file or directory (os error 2)\''
    junk = tmp_path / "junk.fits"
    junk.write_bytes(b"not a fits file".ljust(2880, b" "))
    e = t.call_results(afk.Paths([str(junk)]), 0)[0]
    assert e.status == afk.ERR_IO and "Not a FITS file" in e.message
    # HDU lookup errors (lib.rs:892-906) are raised before any device work
    for args, msg in [((5, ""), "Could not find HDU #5 in FITS file."), ((3, "SCI"), "Could not find HDU 'SCI', nor HDU #3 in FITS file."),
                      ((-1, ""), "'hdu_idx' must be positive, but got -1")]:
        e = t2.call_results(fits, *args)[0]
        assert isinstance(e, afk.AflakError) and e.status == afk.ERR_UNEXPECTED_INPUT and e.message == msg, (args, e.message)
    # and the decode itself has no CPU fallback
    if afk.device_count() == 0:
        e = t2.call_results(fits, 0, "")[0]
        assert isinstance(e, afk.AflakError) and e.status == afk.ERR_CUDA


def test_multi_hdu_layout_and_card_forms(afk, tmp_path):
    a = np.arange(2 * 3 * 5, dtype=">i2").reshape(2, 3, 5)
    b = np.linspace(-1, 1, 7 * 4).astype(">f8").reshape(7, 4)
    blob = (fits_hdu_bytes(None, [("ORIGIN", "it's here"), ("LONGSTR", "x" * 60)]) +
            fits_hdu_bytes(a, [("BSCALE", 0.5), ("BZERO", 32768), ("BLANK", -7), ("BUNIT", "adu")], primary=False, extname="SCI") +
            fits_hdu_bytes(b, [("CRVAL1", 1.5e3), ("CDELT1", -2)], primary=False, extname="ERR"))
    p = tmp_path / "multi.fits"
    p.write_bytes(blob + b"trailing garbage")
    f = afk.Fits.open(p)
    assert len(f) == 3
    assert f.header("NAXIS") == "0" and f.header("ORIGIN") == "it's here" and f.header("LONGSTR") == "x" * 60
    assert f.header("SIMPLE") == "T"
    assert f.header("EXTNAME", 1) == "SCI" and f.header("BITPIX", 1) == "16" and f.header("BLANK", 1) == "-7"
    assert f.header("NAXIS1", 1) == "5" and f.header("NAXIS3", 1) == "2"
    assert f.header("EXTNAME", 2) == "ERR" and f.header("BITPIX", 2) == "-64" and float(f.header("CRVAL1", 2)) == 1500.0
    assert f.header("EXTNAME", 3) is None


def test_oracle_decode_kat(orc):
    # f32: bytes 3F 80 00 00 = 1.0; scaling is always applied: -0.0 * 1 + 0 = +0.0
    raw = bytes([0x3F, 0x80, 0, 0, 0x80, 0, 0, 0, 0xC0, 0x20, 0, 0])
    out = orc.fits_decode(raw, 0, 3)
    assert out.tolist() == [1.0, 0.0, -2.5] and not np.signbit(out[1])
    assert orc.fits_decode(raw, 0, 3, bscale=2.0, bzero=1.0).tolist() == [3.0, 1.0, -4.0]
    # i16 with BLANK and the unsigned-16 convention BZERO = 32768
    raw = np.array([-32768, -1, 0, 32767, 123], dtype=">i2").tobytes()
    out = orc.fits_decode(raw, 2, 5, blank=123, bzero=32768.0)
    assert out[:4].tolist() == [0.0, 32767.0, 32768.0, 65535.0] and np.isnan(out[4])
    # i32: `int as f32` rounds to nearest even above 2^24; two roundings of v*bscale+bzero
    raw = np.array([16777217, -16777219, 2147483647, 5], dtype=">i4").tobytes()
    assert orc.fits_decode(raw, 3, 4).tolist() == [16777216.0, -16777220.0, 2147483648.0, 5.0]
    got = orc.fits_decode(raw, 3, 4, bscale=0.1, bzero=0.3)[3]
    assert got == np.float32(np.float32(5.0) * np.float32(0.1)) + np.float32(0.3)
    # u32 and f64 (`f as f32`)
    assert orc.fits_decode(np.array([4294967295, 7], dtype=">u4").tobytes(), 4, 2).tolist() == [4294967296.0, 7.0]
    d = np.array([0.1, 1e300, -1e-300, np.nan], dtype=">f8")
    out = orc.fits_decode(d.tobytes(), 1, 4)
    assert out[0] == np.float32(0.1) and out[1] == np.inf and out[2] == 0.0 and np.isnan(out[3])
    # numpy agrees on random data for every format
    rng = np.random.default_rng(3)
    for fmt, dt in [(0, ">f4"), (1, ">f8"), (2, ">i2"), (3, ">i4"), (4, ">u4")]:
        a = (rng.normal(0, 1e4, 1000)).astype(dt) if dt[1] == "f" else rng.integers(0, 30000, 1000).astype(dt)
        want = a.astype(np.float32) * np.float32(1.5) + np.float32(-2.0)
        assert np.array_equal(orc.fits_decode(a.tobytes(), fmt, 1000, bscale=1.5, bzero=-2.0), want)


def _raw_header(cards):
    hdr = "".join(cards + ["END".ljust(80)]).encode("ascii")
    return hdr + b" " * (-len(hdr) % 2880)


def test_malformed_headers_are_rejected_not_followed(afk, tmp_path):
    """FITS headers are untrusted input: a product that wraps int64, a negative PCOUNT / GCOUNT below 1 (which made the
    parser walk backwards and re-parse the same HDU for ever) and a data unit larger than the file must end as errors or
    as a short HDU list, never as a tiny buffer under a huge shape."""
    from util import fits_card

    def open_blob(blob, name):
        p = tmp_path / name
        p.write_bytes(blob)
        return afk.get_transform("open_fits").call_results(afk.Paths([str(p)]), 0)[0]

    head = [fits_card("SIMPLE", True), fits_card("BITPIX", -32), fits_card("NAXIS", 3)]
    # NAXIS1*NAXIS2*NAXIS3 = 2^66 wraps to 0 in int64
    e = open_blob(_raw_header(head + [fits_card(f"NAXIS{i}", 1 << 22) for i in (1, 2, 3)]), "wrap.fits")
    assert isinstance(e, afk.AflakError) and "overflows" in e.message
    # (|BITPIX|/8) * count overflows although count itself fits
    e = open_blob(_raw_header(head + [fits_card("NAXIS1", 1 << 31), fits_card("NAXIS2", 1 << 31), fits_card("NAXIS3", 1)]), "wrap2.fits")
    assert isinstance(e, afk.AflakError) and "overflows" in e.message
    ext = [fits_card("XTENSION", "IMAGE"), fits_card("BITPIX", 16), fits_card("NAXIS", 1), fits_card("NAXIS1", 4)]
    prim = fits_hdu_bytes(None, [])
    for pc, gc in ((-2880, 1), (0, 0), (0, -3)):
        e = open_blob(prim + _raw_header(ext + [fits_card("PCOUNT", pc), fits_card("GCOUNT", gc)]) + b"\0" * 2880, f"pg{pc}_{gc}.fits")
        assert isinstance(e, afk.AflakError) and "PCOUNT" in e.message, (pc, gc)
    e = open_blob(_raw_header([fits_card("SIMPLE", True), fits_card("BITPIX", 12), fits_card("NAXIS", 0)]), "bitpix.fits")
    assert isinstance(e, afk.AflakError) and "BITPIX" in e.message
    # a data unit that claims more bytes than the file holds: the header stays readable, the HDU list ends there, and
    # reading the data fails before any allocation (no GPU needed to see the error)
    big = _raw_header(head + [fits_card("NAXIS1", 4096), fits_card("NAXIS2", 4096), fits_card("NAXIS3", 4096)]) + b"\0" * 2880
    f = open_blob(big + fits_hdu_bytes(np.zeros(4, ">f4"), [], primary=False), "truncated.fits")
    assert isinstance(f, afk.Fits) and len(f) == 1 and f.header("NAXIS3") == "4096"
    e = afk.get_transform("fits_to_image").call_results(f, 0, "")[0]
    assert isinstance(e, afk.AflakError) and e.status == afk.ERR_FITS and "end of file" in e.message
