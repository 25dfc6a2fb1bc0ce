"""Shared helpers for the parity tests."""
import numpy as np


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def assert_bit_equal(got, want, what=""):
    """Tier A: bit-exact.  Raw 32-bit patterns are compared on every lane that is not NaN, so -0.0 and +0.0 are different
    results; NaN lanes must coincide (Rust and CUDA do not promise the same NaN payload or sign for 0/0 or Inf-Inf)."""
    got = np.asarray(got, dtype=np.float32)
    want = np.asarray(want, dtype=np.float32)
    assert got.shape == want.shape, f"{what}: shape {got.shape} != {want.shape}"
    gn, wn = np.isnan(got), np.isnan(want)
    assert np.array_equal(gn, wn), f"{what}: NaN positions differ ({gn.sum()} vs {wn.sum()})"
    ok = (bits(got) == bits(want)) | gn
    if not ok.all():
        i = np.argwhere(~ok)[0]
        raise AssertionError(f"{what}: {(~ok).sum()} mismatches, first at {tuple(i)}: got {got[tuple(i)]!r} ({bits(got)[tuple(i)]:#010x}) "
                             f"want {want[tuple(i)]!r} ({bits(want)[tuple(i)]:#010x})")


def ulp_diff(got, want):
    """Distance in units in the last place between two f32 arrays (same-sign finite values)."""
    g = bits(got).astype(np.int64)
    w = bits(want).astype(np.int64)
    g = np.where(g & 0x80000000, 0x80000000 - g, g)
    w = np.where(w & 0x80000000, 0x80000000 - w, w)
    return np.abs(g - w)


def assert_ulp(got, want, max_ulp, what=""):
    """Tier B: within max_ulp of the f32 step-exact oracle; NaN/Inf positions must match exactly."""
    got = np.asarray(got, dtype=np.float32)
    want = np.asarray(want, dtype=np.float32)
    assert got.shape == want.shape
    assert np.array_equal(np.isnan(got), np.isnan(want)), f"{what}: NaN positions differ"
    inf = np.isinf(want)
    assert np.array_equal(got[inf], want[inf]), f"{what}: Inf positions differ"
    fin = np.isfinite(want) & np.isfinite(got)
    assert np.array_equal(np.isfinite(got), np.isfinite(want)), f"{what}: finite positions differ"
    d = ulp_diff(got[fin], want[fin])
    assert d.size == 0 or d.max() <= max_ulp, f"{what}: max ulp diff {d.max()} > {max_ulp}"


def assert_reduction(got, sum64, abs64, what="", divisor=1.0):
    """Tier C: |gpu - o64| <= 1e-5*|o64|  or  <= 4*2^-24*sum|x|; NaN/Inf positions exact."""
    got = np.asarray(got, dtype=np.float32).astype(np.float64)
    o = sum64 / divisor
    a = abs64 / divisor
    assert got.shape == o.shape, f"{what}: shape {got.shape} != {o.shape}"
    assert np.array_equal(np.isnan(got), np.isnan(o)), f"{what}: NaN positions differ"
    inf = np.isinf(o)
    assert np.array_equal(got[inf], o[inf]), f"{what}: Inf positions differ"
    fin = np.isfinite(o)
    err = np.abs(got[fin] - o[fin])
    ok = (err <= 1e-5 * np.abs(o[fin])) | (err <= 4 * 2.0 ** -24 * a[fin])
    assert ok.all(), f"{what}: {(~ok).sum()} pixels outside tolerance, worst rel err {np.max(err / np.maximum(np.abs(o[fin]), 1e-300))}"


# ---- a minimal FITS writer for the ingest tests (FITS standard 4.0: 80-char cards, 2880-byte blocks, big-endian data) ----
def fits_card(key, value=None, comment=""):
    if value is None:
        return f"{key:<8}".ljust(80)
    if isinstance(value, bool):
        v = f"{'T' if value else 'F':>20}"
    elif isinstance(value, (int, np.integer)):
        v = f"{int(value):>20}"
    elif isinstance(value, (float, np.floating)):
        v = f"{float(value):>20.12E}"
    else:
        v = "'" + str(value).replace("'", "''").ljust(8) + "'"
    c = f"{key:<8}= {v}"
    if comment:
        c += f" / {comment}"
    assert len(c) <= 80, c
    return c.ljust(80)


def fits_hdu_bytes(data_be, cards, primary=True, extname=None):
    """data_be: numpy array in its on-disk big-endian dtype (or None); cards: list of (key, value) after NAXISn."""
    bitpix = {">f4": -32, ">f8": -64, ">i2": 16, ">i4": 32, ">u4": 32, "|u1": 8, ">u1": 8}
    out = [fits_card("SIMPLE", True) if primary else fits_card("XTENSION", "IMAGE")]
    if data_be is None:
        out += [fits_card("BITPIX", 8), fits_card("NAXIS", 0)]
    else:
        out += [fits_card("BITPIX", bitpix[data_be.dtype.str]), fits_card("NAXIS", data_be.ndim)]
        out += [fits_card(f"NAXIS{i + 1}", n) for i, n in enumerate(reversed(data_be.shape))]
    if not primary:
        out += [fits_card("PCOUNT", 0), fits_card("GCOUNT", 1)]
    if extname is not None:
        out.append(fits_card("EXTNAME", extname))
    out += [fits_card(k, v) for k, v in cards]
    out.append("END".ljust(80))
    hdr = "".join(out).encode("ascii")
    hdr += b" " * (-len(hdr) % 2880)
    data = b"" if data_be is None else data_be.tobytes()
    data += b"\0" * (-len(data) % 2880)
    return hdr + data


def build_c_client(tmpdir):
    """Compile tests/c/abi_client.c (plain C99, only include/aflak_b200.h and the shared library) and return the binary's path."""
    import os
    import subprocess

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = os.path.join(str(tmpdir), "abi_client")
    env = dict(os.environ)
    env.pop("CC", None)
    subprocess.check_call(["gcc", "-std=c99", "-O2", "-Wall", "-Wextra", "-Werror", "-I", os.path.join(root, "include"),
                           os.path.join(root, "tests", "c", "abi_client.c"), "-L", os.path.join(root, "aflak_b200"), "-laflak_b200",
                           "-Wl,-rpath," + os.path.join(root, "aflak_b200"), "-lm", "-o", exe], env=env)
    return exe
