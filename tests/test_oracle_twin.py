"""CPU: harden the (unpinned) oracle as far as this box allows.

1. The reference's own data fixture, whole: tests/golden/test.fits is aflak_primitives/test/test.fits (1.9 MB, BITPIX -32,
   200 x 49 x 49, reference-held data, not source).  The oracle's byte-by-byte FITS decode must reproduce numpy's
   big-endian read of the data unit bit for bit, and the header facts committed in test_fits_facts.json.
2. An independently written numpy-f32 twin of the axis-0 reductions (tests/twin.py, no code shared with
   oracle/aflak_oracle.c): twin == C oracle bit for bit on that cube, on seeded synthetic cubes and on adversarial values
   (signed zeros, denormals, infinities, cancellation).  One typo cannot sit in both restatements.
"""
import json
import os

import numpy as np
import pytest

import twin
from util import assert_bit_equal

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def reference_cube():
    raw = open(os.path.join(GOLD, "test.fits"), "rb").read()
    # primary header: one 2880-byte block (36 cards, END inside), data unit big-endian f32 [200][49][49], file not padded
    assert raw[:30].startswith(b"SIMPLE  =") and b"END" + b" " * 77 in raw[:2880]
    data = raw[2880:2880 + 200 * 49 * 49 * 4]
    return data, np.frombuffer(data, dtype=">f4").astype(np.float32).reshape(200, 49, 49)


def test_reference_fixture_decodes_bit_for_bit(orc):
    data, want = reference_cube()
    got = orc.fits_decode(data, 0, want.size).reshape(want.shape)
    assert_bit_equal(got, want, "oracle FITS decode of the reference's test.fits")
    facts = json.load(open(os.path.join(GOLD, "test_fits_facts.json")))
    assert list(want.shape) == facts["shape"] and int(np.isnan(want).sum()) == facts["nan_count"]
    assert float(want.max()) == facts["data_max"] and float(want.min()) == facts["data_min"]
    # header DATAMAX / DATAMIN are the 7-digit decimal forms of the data's extremes
    assert abs(facts["datamax_header"] - float(want.max())) < 1e-5 and abs(facts["datamin_header"] - float(want.min())) < 1e-6
    crop = np.load(os.path.join(GOLD, "test_fits_crop.npy"))
    assert_bit_equal(want[:, 18:30, 18:30], crop, "committed crop is the same data")


def cubes():
    _, ref = reference_cube()
    yield "test.fits", ref, [(0, 200), (40, 160), (7, 8), (199, 200), (10, 13)]
    rng = np.random.default_rng(0xAF1A0080)
    a = (rng.standard_normal((37, 5, 9)) * 100).astype(np.float32)
    yield "normal", a, [(0, 37), (3, 20), (36, 37)]
    b = np.exp(rng.standard_normal((64, 4, 6)) * 8).astype(np.float32)  # 30 decades of dynamic range: every rounding matters
    b[::7] *= -1
    yield "wide range", b, [(0, 64), (5, 60)]
    c = rng.standard_normal((24, 3, 4)).astype(np.float32)
    c[3, 0, 0] = np.inf
    c[5, 0, 1] = -np.inf
    c[4, 0, 2] = np.inf
    c[9, 0, 2] = -np.inf       # inf + -inf -> NaN in the sums
    c[:, 1, 0] = 0.0
    c[::2, 1, 0] = -0.0        # signed zeros only
    c[:, 1, 1] = -0.0
    c[:, 1, 2] = 1e-42         # denormals
    c[:, 2, 0] = 16777216.0    # 2^24: +1 is lost
    c[1::2, 2, 0] = 1.0
    yield "special values", c, [(0, 24), (2, 11)]


@pytest.mark.parametrize("name,cube,windows", list(cubes()), ids=[c[0] for c in cubes()])
def test_twin_equals_oracle_sum_mean(orc, name, cube, windows):
    with np.errstate(all="ignore"):
        for s, e in windows:
            assert_bit_equal(orc.integral(cube, s, e), twin.sum_axis(cube, s, e), f"{name} sum_axis [{s},{e})")
            assert_bit_equal(orc.average(cube, s, e), twin.mean_axis(cube, s, e), f"{name} mean_axis [{s},{e})")


@pytest.mark.parametrize("name,cube,windows", list(cubes()), ids=[c[0] for c in cubes()])
def test_twin_equals_oracle_variance_stddev(orc, name, cube, windows):
    with np.errstate(all="ignore"):
        for s, e in windows:
            if e - s < 2:  # ndarray 0.12.1 var_axis panics: ddof = 1 is not smaller than the axis length
                with pytest.raises(Exception):
                    orc.variance(cube, s, e)
                continue
            assert_bit_equal(orc.variance(cube, s, e), twin.var_axis(cube, s, e), f"{name} var_axis [{s},{e})")
            assert_bit_equal(orc.stddev(cube, s, e), twin.std_axis(cube, s, e), f"{name} std_axis [{s},{e})")


@pytest.mark.parametrize("name,cube,windows", list(cubes())[:3], ids=[c[0] for c in list(cubes())[:3]])
def test_twin_equals_oracle_median_fold_argmap_centroid(orc, name, cube, windows):
    with np.errstate(all="ignore"):
        for s, e in windows:
            assert_bit_equal(orc.median(cube, s, e), twin.median(cube, s, e), f"{name} median [{s},{e})")
            for is_min in (False, True):
                assert_bit_equal(orc.minmax_fold(cube, s, e, is_min), twin.fold_minmax(cube, s, e, is_min), f"{name} fold {is_min}")
                for rng in (0, 3, -2):
                    assert_bit_equal(orc.argmap(cube, s, e, rng, is_min), twin.argmap(cube, s, e, rng, is_min), f"{name} argmap rng={rng} min={is_min}")
            assert_bit_equal(orc.centroid(cube, s, e), twin.centroid(cube, s, e), f"{name} centroid [{s},{e})")


def test_twin_median_signed_zeros_and_ties(orc):
    # slice::sort_by(partial_cmp) is stable: equal keys (-0.0 == 0.0) keep their frame order
    col = np.array([0.0, -0.0, 1.0, -1.0, -0.0, 0.0], np.float32).reshape(6, 1, 1)
    for s, e in [(0, 6), (0, 5), (1, 5), (0, 2), (1, 2)]:
        assert_bit_equal(orc.median(col, s, e), twin.median(col, s, e), f"median [{s},{e})")


def test_twin_fma_is_a_single_rounding():
    # a*b + c where the double rounding of (f64 sum -> f32) would differ from the fused result
    a = np.float32(1.0 + 2.0 ** -12)
    b = np.float32(1.0 + 2.0 ** -12)
    c = np.float32(2.0 ** -25)
    exact = float(a) * float(b) + float(c)  # exact in f64? the product needs 25 bits, the sum fits 53: yes
    assert twin.fma_f32(a, b, c) == np.float32(exact)
    # halfway case: p + c lies exactly between two f32 values in f64 only after rounding; round-to-odd keeps the sticky bit
    x = np.float32(1.0 + 2.0 ** -23)
    y = np.float32(1.0 - 2.0 ** -24)
    z = np.float32(2.0 ** -60)
    from fractions import Fraction
    want = Fraction(float(x)) * Fraction(float(y)) + Fraction(float(z))
    lo = np.float32(float(want))
    cand = [np.nextafter(lo, np.float32(-np.inf)), lo, np.nextafter(lo, np.float32(np.inf))]
    best = min(cand, key=lambda v: abs(Fraction(float(v)) - want))
    assert twin.fma_f32(x, y, z) == best
