#!/usr/bin/env python
"""Regenerates the committed fixtures under tests/golden/ (run in the build container, where /root/reference exists).

Three kinds of fixture, none of which is produced by code of this repository:

  lut_golden.json        the value-asserting vectors of the reference's own unit tests for the colour LUT
                         (aflak_plot/src/imshow/lut.rs:292-343): 11 colour triples and the 4 bounds() tuples,
                         transcribed as data.
  test_fits_crop.npy     a [200, 12, 12] f32 crop (rows/cols 18..30) of the reference's only data fixture,
  test_fits_facts.json   aflak_primitives/test/test.fits (49x49x200, BITPIX -32, BUNIT K), decoded here by a
                         20-line FITS reader (big-endian f32, BSCALE 1 / BZERO 0), plus the header facts and
                         f64 column sums of the crop computed with numpy -- a real-data input for the parity tests.
  test_fits_crop.fits    the same crop written back as a FITS file under the reference file's own header cards
                         (NAXIS1/2 and CRPIX1/2 adjusted to the crop): input of the header-parser / ingest tests.
  kat.json               hand-derivable known answers for the operators that have no value-asserting reference test
                         (SURVEY.md 8c): closed forms worked out from the reference's formulas, and the error
                         strings of aflak_primitives/src/precond.rs.

The reference itself cannot be executed here (pure Rust, no rustc/cargo in the image), so there are no
reference-generated output vectors beyond these.
"""
import json
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/src"


def lut_golden():
    g1 = [[0.0, [0, 0, 255]], [0.5, [255, 255, 255]], [1.0, [255, 0, 0]]]
    g2 = [[0.0, [0, 0, 0]], [1.0, [255, 255, 255]]]
    return {
        "source": "aflak_plot/src/imshow/lut.rs:292-343",
        "color_at": [
            {"gradient": g1, "lims": [0.0, 0.5, 1.0], "cases": [[0.0, [0, 0, 255]], [1.0, [255, 0, 0]], [0.5, [254, 254, 255]], [0.25, [127, 127, 255]]]},
            # lims assigned after construction: the table was built with the default lims (lut.rs:332-334)
            {"gradient": g2, "lims": [0.2, 0.5, 0.9], "table_lims": [0.0, 0.5, 1.0],
             "cases": [[0.0, [0, 0, 0]], [0.1, [0, 0, 0]], [0.2, [0, 0, 0]], [0.55, [127, 127, 127]], [0.9, [255, 255, 255]], [0.95, [255, 255, 255]], [1.0, [255, 255, 255]]]},
        ],
        "bounds": {"gradient": g1, "lims": [0.0, 0.5, 1.0],
                   "pairs": [[[0.0, [0, 0, 255]], [0.0, [0, 0, 255]]], [[0.0, [0, 0, 255]], [0.5, [255, 255, 255]]],
                             [[0.5, [255, 255, 255]], [1.0, [255, 0, 0]]], [[1.0, [255, 0, 0]], [1.0, [255, 0, 0]]]]},
    }


def read_fits(path):
    b = open(path, "rb").read()
    hdr, i = {}, 0
    while True:
        card = b[i:i + 80].decode("ascii")
        i += 80
        if card.startswith("END"):
            break
        if "=" in card[:10]:
            k, v = card.split("=", 1)
            hdr[k.strip()] = v.split("/")[0].strip().strip("'").strip()
    off = (i + 2879) // 2880 * 2880
    shape = [int(hdr[f"NAXIS{a}"]) for a in range(int(hdr["NAXIS"]), 0, -1)]  # [NAXISn..NAXIS1], fits.rs:42-49
    assert int(hdr["BITPIX"]) == -32
    data = np.frombuffer(b, dtype=">f4", count=int(np.prod(shape)), offset=off).reshape(shape).astype(np.float32)
    return hdr, data


def fits_crop():
    hdr, data = read_fits(os.path.join(REF, "aflak_primitives/test/test.fits"))
    crop = np.ascontiguousarray(data[:, 18:30, 18:30])
    np.save(os.path.join(HERE, "test_fits_crop.npy"), crop)
    facts = {
        "source": "aflak_primitives/test/test.fits", "shape": list(data.shape), "bitpix": int(hdr["BITPIX"]), "bunit": hdr["BUNIT"],
        "bscale": float(hdr["BSCALE"]), "bzero": float(hdr["BZERO"]), "ctype": [hdr["CTYPE1"], hdr["CTYPE2"], hdr["CTYPE3"]],
        "datamax_header": float(hdr["DATAMAX"]), "datamin_header": float(hdr["DATAMIN"]),
        "data_max": float(np.nanmax(data)), "data_min": float(np.nanmin(data)), "nan_count": int(np.isnan(data).sum()),
        "crop": "data[:, 18:30, 18:30]",
        # numpy f64 sums of the crop over [40,160): an oracle-independent check value for integral/average
        "crop_sum_f64_40_160": crop[40:160].astype(np.float64).sum(0).tolist(),
        "crop_frame_100": crop[100].astype(np.float64).tolist(),
    }
    return facts


def fits_crop_file():
    """test_fits_crop.fits: the same crop as a FITS file -- the reference file's own header cards (free-format values and
    all), with NAXIS1/NAXIS2 set to the crop size and CRPIX1/CRPIX2 moved by the crop origin, followed by the crop as
    big-endian f32.  Input of the header-parser / ingest tests (the full 1.9 MB file stays in /root/reference)."""
    b = open(os.path.join(REF, "aflak_primitives/test/test.fits"), "rb").read()
    cards, i = [], 0
    while True:
        card = b[i:i + 80].decode("ascii")
        i += 80
        cards.append(card)
        if card.startswith("END"):
            break
    out = []
    for c in cards:
        k = c[:8].strip()
        if k in ("NAXIS1", "NAXIS2"):
            c = f"{k:<8}=      12 /".ljust(80)
        elif k in ("CRPIX1", "CRPIX2"):
            v = float(c.split("=", 1)[1].split("/")[0])
            c = f"{k:<8}=      {v - 18.0:.6E} /".ljust(80)
        out.append(c)
    hdr = "".join(out).encode("ascii")
    hdr += b" " * (-len(hdr) % 2880)
    crop = np.load(os.path.join(HERE, "test_fits_crop.npy"))
    data = crop.astype(">f4").tobytes()
    data += b"\0" * (-len(data) % 2880)
    with open(os.path.join(HERE, "test_fits_crop.fits"), "wb") as f:
        f.write(hdr + data)


def kat():
    f32 = np.float32
    return {
        "source": "closed forms from aflak_primitives/src/lib.rs formulas; messages from precond.rs:3-114",
        # cube[k,y,x] = k: integral over [a,b) = (a+b-1)(b-a)/2 exactly (all partial sums < 2^24)
        "integral_ramp": [{"nz": 64, "a": a, "b": b, "value": (a + b - 1) * (b - a) / 2} for a, b in [(0, 64), (3, 17), (10, 11), (63, 64)]],
        "middle_width": [{"a": 3134, "b": 3154, "middle": 3144.0, "width": 20.0}, {"a": 0, "b": 1, "middle": 0.5, "width": 1.0}],
        # create_equivalent_width: ((off-on)*fl/off)*(-1 if emission) ; > max -> 0   (lib.rs:1584-1596)
        "equivalent_width": [
            {"off": 2.0, "on": 1.0, "fl": 20.0, "max": 1e10, "emission": True, "value": -10.0},
            {"off": 2.0, "on": 1.0, "fl": 20.0, "max": 1e10, "emission": False, "value": 10.0},
            {"off": 2.0, "on": 1.0, "fl": 20.0, "max": 5.0, "emission": False, "value": 0.0},
            {"off": 4.0, "on": 8.0, "fl": 3.0, "max": 1e10, "emission": True, "value": 3.0},
        ],
        "ratio_from_bands": {"z": 3144.0, "z1": 3109.0, "z2": 3184.0,
                             "off_ratio_1": float(f32(f32(35.0) / f32(75.0))), "off_ratio_2": float(f32(1.0) - f32(f32(35.0) / f32(75.0)))},
        # down_sampling n=3 on a 4x5 image keeps (0,0),(0,3),(3,0),(3,3) (lib.rs:1894-1962)
        "down_sampling": {"shape": [4, 5], "n": 3, "out_shape": [2, 2], "picks": [[0, 0], [0, 3], [3, 0], [3, 3]]},
        # apply_tone_curve with the default 256-entry identity table (lib.rs:1858-1871): key = trunc(v*255/65535)
        "tone_curve_default": [{"v": 65535.0, "out": 65535.0}, {"v": 257.0, "out": 257.0}, {"v": 0.0, "out": 0.0},
                               {"v": 300.0, "out": 257.0}, {"v": -5.0, "out": 0.0}],
        # convert_to_logscale(a=1000): v=vmin -> ln(1)/ln(a) = 0
        "logscale": [{"v": 2.0, "a": 1000.0, "vmin": 2.0, "vmax": 9.0, "out": 0.0}],
        "linear_composition": [{"u": 1.5, "v": -2.0, "a": 2.0, "b": 0.25, "out": 2.5}],
        "image_min_max_empty": {"min": float(np.finfo(np.float32).max), "max": float(-np.finfo(np.float32).max)},
        "errors": {
            "start_negative": "'start' must be positive, but got -1",
            "end_negative": "'end' must be positive, but got -3",
            "start_ge_end": "Expected start < end, but got 5 < 3",
            "end_too_large_nz10": "'end' greater or equal to the input image 'image''s frame count (expected 11 < 11)",
            "frame_too_large_nz10": "'frame_idx' greater or equal to the input image 'input_img''s frame count (expected 10 < 10)",
            "down_sampling_n0": "n should be greater than 0, 0 > 0",
            "lincomp_shape": "Cannot compose images of different dimensions ('i1' has dimension [2, 3], while 'i2' has dimension [3, 2])",
            "ratio_from_bands_order": "wrong magnitude correlation (5 < 3 < 9)",
        },
    }


if __name__ == "__main__":
    json.dump(lut_golden(), open(os.path.join(HERE, "lut_golden.json"), "w"), indent=1)
    json.dump(kat(), open(os.path.join(HERE, "kat.json"), "w"), indent=1)
    if os.path.exists(REF):
        json.dump(fits_crop(), open(os.path.join(HERE, "test_fits_facts.json"), "w"), indent=1)
        fits_crop_file()
    print(sorted(os.listdir(HERE)))
