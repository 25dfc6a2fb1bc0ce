"""A second, independently written restatement of the reference's axis-0 reductions: numpy float32, one explicit step per
frame, vectorised over pixels only.  Test infrastructure: it shares no code with oracle/aflak_oracle.c, so that a slip in
one restatement cannot hide in the other (tests/test_oracle_golden.py asserts twin == C oracle bit for bit).

What is restated (aflak-vis/aflak, src/aflak_primitives/src/lib.rs and ndarray 0.12.1, Cargo.lock:1112):
  sum_axis     lib.rs:1324-1326 -> ndarray sum_axis(Axis(0)): res = frame[0]; res = res + frame[k], k = 1..n, in f32
  mean_axis    lib.rs:1328-1330 -> sum_axis / cnt, cnt built by n-1 additions of 1.0f32
  var_axis     lib.rs:1332-1338 -> ndarray var_axis(Axis(0), ddof = 1.0): Welford, the update of sum_sq is one mul_add (fused)
  median       lib.rs:1340-1372  stable sort by partial_cmp, middle element or the f32 mean of the two middle ones
  fold min/max lib.rs:1374-1395  `if x > y { x } else { y }` folds starting at -inf / +inf
  argmap       lib.rs:1397-1459  strict comparison: the first extremum wins; out = (k + start + range) as f32 (no WCS)
  centroid     lib.rs:1466-1527  two sequential f32 sums (flux, flux * wavelength rounded separately), then a true division
"""
import numpy as np

F = np.float32


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def sum_axis(cube, start, end):
    cube = _f32(cube)
    res = cube[start].copy()
    for k in range(start + 1, end):
        res = (res + cube[k]).astype(F)
    return res


def count_f32(n):
    c = F(1.0)
    for _ in range(1, n):
        c = F(c + F(1.0))
    return c


def mean_axis(cube, start, end):
    with np.errstate(all="ignore"):
        return (sum_axis(cube, start, end) / count_f32(end - start)).astype(F)


def fma_f32(a, b, c):
    """round_f32(a*b + c) with a single rounding.  a*b is exact in f64 (24 + 24 bits); the f64 sum is made round-to-odd
    with the TwoSum error term, which makes the final f64 -> f32 rounding the correct one (f64 keeps 29 spare bits)."""
    a, b, c = (np.asarray(x, dtype=np.float64) for x in (a, b, c))
    with np.errstate(all="ignore"):
        p = a * b
        s = p + c
        bb = s - p
        err = (p - (s - bb)) + (c - bb)
        fin = np.isfinite(s) & np.isfinite(err) & (err != 0)
        even = (s.view(np.int64) & 1) == 0
        toward = np.where(err > 0, np.inf, -np.inf)
        s = np.where(fin & even, np.nextafter(s, toward), s)
        return s.astype(F)


def var_axis(cube, start, end, ddof=1.0):
    cube = _f32(cube)
    mean = np.zeros(cube.shape[1:], F)
    sum_sq = np.zeros(cube.shape[1:], F)
    count = F(0)
    with np.errstate(all="ignore"):
        for i, k in enumerate(range(start, end)):
            count = F(i + 1)
            x = cube[k]
            delta = (x - mean).astype(F)
            mean = (mean + (delta / count).astype(F)).astype(F)
            sum_sq = fma_f32((x - mean).astype(F), delta, sum_sq)
        return (sum_sq / F(count - F(ddof))).astype(F)


def std_axis(cube, start, end, ddof=1.0):
    with np.errstate(all="ignore"):
        return np.sqrt(var_axis(cube, start, end, ddof)).astype(F)


def median(cube, start, end):
    cube = _f32(cube)
    v = np.sort(cube[start:end], axis=0, kind="stable")
    n = end - start
    if n % 2 == 1:
        return v[(n - 1) // 2].copy()
    with np.errstate(all="ignore"):
        return ((v[n // 2] + v[n // 2 - 1]).astype(F) / F(2.0)).astype(F)


def fold_minmax(cube, start, end, is_min):
    cube = _f32(cube)
    acc = np.full(cube.shape[1:], np.inf if is_min else -np.inf, F)
    for k in range(start, end):
        y = cube[k]
        take_acc = (acc < y) if is_min else (acc > y)
        acc = np.where(take_acc, acc, y)
    return acc


def argmap(cube, start, end, rng, is_min):
    cube = _f32(cube)
    value = np.full(cube.shape[1:], np.inf if is_min else -np.inf, F)
    out = np.zeros(cube.shape[1:], F)
    for k in range(start, end):
        x = cube[k]
        hit = (x < value) if is_min else (x > value)
        value = np.where(hit, x, value)
        out = np.where(hit, F(k + rng), out)
    return out


def centroid(cube, start, end, world=None):
    """world: optional callable pixel(f32) -> f32 (pix2world); None = the index itself (lib.rs:1507-1510)."""
    cube = _f32(cube)
    flux_sum = np.zeros(cube.shape[1:], F)
    wave = np.zeros(cube.shape[1:], F)
    with np.errstate(all="ignore"):
        for k in range(start, end):
            flux = cube[k]
            w = F(k) if world is None else F(world(F(k)))
            flux_sum = (flux_sum + flux).astype(F)
            wave = (wave + (flux * w).astype(F)).astype(F)
        return (wave / flux_sum).astype(F)
