"""GPU parity: chained operators through the mini evaluator (device-handle cache), the fused kernels it substitutes,
and the fused-gather reduce path.

Reference: src/src/templates.rs:143-463 (equivalent-width template), aflak_cake/src/dst/compute.rs:153-274 and
cache.rs:18-148 (memoised evaluation), aflak_cake/tests/dst.rs:114-143 (cache invalidation).
The unfused operator-by-operator path is the parity reference for every fused kernel: in SUM_SEQUENTIAL mode both must
be bit-identical to the oracle chain; in the default compensated mode the fused and unfused device paths agree to the
last bit or two (their launches split the spectral axis between CTAs differently), and with the f64 oracle within the
Tier C tolerance.
"""
import numpy as np
import pytest

from util import assert_bit_equal, assert_reduction, assert_ulp

pytestmark = pytest.mark.gpu

ON, OFF1, OFF2 = (134, 154), (99, 119), (174, 194)


def build_ew_graph(g, cube_const, windows=(ON, OFF1, OFF2), fl=20.0, vmax=1e10, emission=True):
    """The DAG of the equivalent_width template, node for node."""
    avg = []
    for s, e in windows:
        t = g.add_transform("average")
        g.connect(cube_const, 0, t, 0)
        g.connect(g.add_constant(int(s)), 0, t, 1)
        g.connect(g.add_constant(int(e)), 0, t, 2)
        avg.append(t)
    on, off1, off2 = avg
    rb = g.add_transform("ratio_from_bands")
    g.connect(on, 1, rb, 0)
    g.connect(off1, 1, rb, 1)
    g.connect(off2, 1, rb, 2)
    lc = g.add_transform("linear_composition")
    g.connect(off1, 0, lc, 0)
    g.connect(off2, 0, lc, 1)
    g.connect(rb, 0, lc, 2)
    g.connect(rb, 1, lc, 3)
    ew = g.add_transform("create_equivalent_width")
    g.connect(lc, 0, ew, 0)
    g.connect(on, 0, ew, 1)
    g.connect(g.add_constant(float(fl)), 0, ew, 2)
    g.connect(g.add_constant(float(vmax)), 0, ew, 3)
    g.connect(g.add_constant(bool(emission)), 0, ew, 4)
    mm = g.add_transform("image_min_max")
    g.connect(ew, 0, mm, 0)
    ls = g.add_transform("convert_to_logscale")
    g.connect(ew, 0, ls, 0)
    g.connect(g.add_constant(1000.0), 0, ls, 1)
    g.connect(mm, 0, ls, 2)
    g.connect(mm, 1, ls, 3)
    return {"avg": avg, "rb": rb, "lc": lc, "ew": ew, "mm": mm, "ls": ls}


def oracle_ew(orc, host, windows=(ON, OFF1, OFF2), fl=20.0, vmax=1e10, emission=True):
    on, off1, off2 = [orc.average(host, s, e) for s, e in windows]
    mid = [orc.middle_width(s, e)[0] for s, e in windows]
    r2, r1 = orc.ratio_from_bands(mid[0], mid[1], mid[2])
    off = orc.linear_composition(off1, off2, r2, r1)
    return orc.equivalent_width(off, on, fl, vmax, emission)


@pytest.mark.parametrize("shape", [(200, 74, 74), (200, 49, 49), (200, 16, 300)])
def test_ew_template_sequential_bit_exact(gpu, orc, shape):
    nz, ny, nx = shape
    host = orc.gen_cube(31, nz, ny, nx)
    want = oracle_ew(orc, host)
    cube = gpu.Image.upload(host)
    cube.unit = "K"
    gpu.set_sum_mode(gpu.SUM_SEQUENTIAL)
    try:
        results = {}
        for fusion in (True, False):
            g = gpu.Graph()
            g.set_fusion(fusion)
            n = build_ew_graph(g, g.add_constant(cube))
            ew = g.compute(n["ew"])
            results[fusion] = ew.download()
            assert_bit_equal(results[fusion], want, f"EW template fusion={fusion}")
            assert ew.unit == "Ang" and ew.tag is None
            mn, mx = g.compute(n["mm"], 0), g.compute(n["mm"], 1)
            assert (mn, mx) == orc.image_min_max(want)
            ls = g.compute(n["ls"]).download()
            ols = orc.convert_to_logscale(want, 1000.0, mn, mx)
            assert np.nanmax(np.abs(ls.astype(np.float64) - ols)) <= 4 * 2.0 ** -24 * max(1.0, np.nanmax(np.abs(ols)))
            # fused: one kernel for the five reduction/elementwise nodes; unfused: 3 averages + rb + lc + ew
            assert g.calls == (3 if fusion else 8), g.calls
        assert_bit_equal(results[True], results[False])
    finally:
        gpu.set_sum_mode(gpu.SUM_COMPENSATED)


def test_ew_template_compensated(gpu, orc):
    nz, ny, nx = 300, 74, 74
    host = orc.gen_cube(32, nz, ny, nx)
    cube = gpu.Image.upload(host)
    windows = ((100, 200), (20, 90), (210, 290))
    out = {}
    for fusion in (True, False):
        g = gpu.Graph()
        g.set_fusion(fusion)
        n = build_ew_graph(g, g.add_constant(cube), windows)
        out[fusion] = g.compute(n["ew"]).download()
    # the fused and unfused launches split the spectral axis differently, so the compensated means may differ in the
    # last bit; EW amplifies that by off/(off-on) -- compare at that scale (bit-exactness is the sequential-mode test)
    np.testing.assert_allclose(out[True], out[False], rtol=1e-4, atol=1e-4)
    # against exact (f64) window means pushed through the f32 operator chain: the result is a quotient of O(1)
    # quantities, so a few f32 ulps of slack on the means translate to ~1e-6 relative on EW*off/fl
    m = [orc.sum_f64(host, s, e)[0] / (e - s) for s, e in windows]
    mid = [orc.middle_width(s, e)[0] for s, e in windows]
    r2, r1 = orc.ratio_from_bands(*[mid[0], mid[1], mid[2]])
    off = np.float64(r2) * m[1] + np.float64(r1) * m[2]
    want = -((off - m[0]) * 20.0 / off)
    np.testing.assert_allclose(out[True], want, rtol=1e-5, atol=2e-5)


def test_fused_ew_entry_point_and_minmax(gpu, orc):
    host = orc.gen_cube(33, 200, 40, 44, pathology=True)
    cube = gpu.Image.upload(host)
    gpu.set_sum_mode(gpu.SUM_SEQUENTIAL)
    try:
        img, (mn, mx) = gpu.fused_equivalent_width(cube, ON, OFF1, OFF2, 20.0, 1e10, True, want_minmax=True)
        want = oracle_ew(orc, host)
        assert_bit_equal(img.download(), want)
        omn, omx = orc.image_min_max(want)
        assert (mn, mx) == (omn, omx)
    finally:
        gpu.set_sum_mode(gpu.SUM_COMPENSATED)
    with pytest.raises(gpu.AflakError) as ei:
        gpu.fused_equivalent_width(cube, ON, OFF2, OFF1, 20.0, 1e10, True)  # band centres out of order
    assert "wrong magnitude correlation" in ei.value.message
    with pytest.raises(gpu.AflakError) as ei:
        gpu.fused_equivalent_width(cube, (190, 201), OFF1, OFF2, 20.0, 1e10, True)
    assert "frame count" in ei.value.message


def test_fused_logscale_minmax_matches_the_two_operators(gpu, orc):
    """convert_to_logscale fed by image_min_max of the same image: the fused path keeps the range on the device; same bits as
    the two operators, including images with NaN / Inf, a constant image (zero range) and an all-NaN image (the operator's
    FLT_MAX / -FLT_MAX initial values)."""
    rng = np.random.default_rng(0xAF1A0061)
    base = np.exp(rng.standard_normal((67, 129))).astype(np.float32)
    holes = base.copy()
    holes[::7, ::5] = np.nan
    holes[3, 4] = np.inf
    holes[5, 6] = -np.inf
    for tag, host in [("lognormal", base), ("nan/inf", holes), ("constant", np.full((8, 12), 2.5, np.float32)),
                      ("all nan", np.full((4, 8), np.nan, np.float32)), ("negative", -base)]:
        img = gpu.Image.upload(host)
        mn, mx = gpu.call("image_min_max", img)
        want = gpu.call("convert_to_logscale", img, 1000.0, float(mn), float(mx)).download()
        got = gpu.fused_logscale_minmax(img, 1000.0).download()
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32)) or (
            np.array_equal(np.isnan(got), np.isnan(want)) and np.array_equal(got[~np.isnan(got)], want[~np.isnan(want)])), tag
    # through the evaluator: the pattern is substituted when fusion is on, and the results agree with fusion off
    outs = []
    for fusion in (True, False):
        g = gpu.Graph()
        g.set_fusion(fusion)
        c = g.add_constant(gpu.Image.upload(holes))
        mm = g.add_transform("image_min_max")
        g.connect(c, 0, mm, 0)
        ls = g.add_transform("convert_to_logscale")
        g.connect(c, 0, ls, 0)
        g.connect(g.add_constant(1000.0), 0, ls, 1)
        g.connect(mm, 0, ls, 2)
        g.connect(mm, 1, ls, 3)
        outs.append((g.compute(ls).download(), g.calls))
    assert np.array_equal(np.isnan(outs[0][0]), np.isnan(outs[1][0]))
    assert np.array_equal(outs[0][0][~np.isnan(outs[0][0])], outs[1][0][~np.isnan(outs[1][0])])
    assert outs[0][1] == 1 and outs[1][1] == 2  # fused: one call; unfused: image_min_max + convert_to_logscale


def test_average_ratio_graph_fused_vs_unfused(gpu, orc):
    """BASELINE config 3 on a small cube: average, average, ratio."""
    nz, ny, nx = 128, 64, 96
    host = orc.gen_cube(34, nz, ny, nx)
    cube = gpu.Image.upload(host)
    cube.unit = "Jy/beam"
    res = {}
    for mode in (gpu.SUM_SEQUENTIAL, gpu.SUM_COMPENSATED):
        gpu.set_sum_mode(mode)
        try:
            for fusion in (True, False):
                g = gpu.Graph()
                g.set_fusion(fusion)
                c = g.add_constant(cube)
                a, b = g.add_transform("average"), g.add_transform("average")
                for t, (s, e) in ((a, (0, nz // 2)), (b, (nz // 2, nz))):
                    g.connect(c, 0, t, 0)
                    g.connect(g.add_constant(s), 0, t, 1)
                    g.connect(g.add_constant(e), 0, t, 2)
                r = g.add_transform("ratio")
                g.connect(a, 0, r, 0)
                g.connect(b, 0, r, 1)
                out = g.compute(r)
                res[(mode, fusion)] = out.download()
                assert out.unit == "" and out.shape == (ny, nx)
                assert g.calls == (1 if fusion else 3)
        finally:
            gpu.set_sum_mode(gpu.SUM_COMPENSATED)
    want = orc.ratio(orc.average(host, 0, nz // 2), orc.average(host, nz // 2, nz))
    assert_bit_equal(res[(gpu.SUM_SEQUENTIAL, True)], want)
    assert_bit_equal(res[(gpu.SUM_SEQUENTIAL, False)], want)
    assert_ulp(res[(gpu.SUM_COMPENSATED, True)], res[(gpu.SUM_COMPENSATED, False)], 4, "fused vs unfused, compensated")
    s0, _ = orc.sum_f64(host, 0, nz // 2)
    s1, _ = orc.sum_f64(host, nz // 2, nz)
    np.testing.assert_allclose(res[(gpu.SUM_COMPENSATED, True)], s0 / s1, rtol=1e-6)


def test_cache_holds_device_handles_and_invalidates(gpu, orc):
    """Port of aflak_cake/tests/dst.rs:114-143 onto device images: a cached output is the same device buffer until an
    ancestor changes; dragging the range (update_constant_node) recomputes only the dependants."""
    host = orc.gen_cube(35, 64, 20, 24)
    cube = gpu.Image.upload(host)
    g = gpu.Graph()
    c = g.add_constant(cube)
    s, e = g.add_constant(0), g.add_constant(32)
    avg = g.add_transform("average")
    g.connect(c, 0, avg, 0)
    g.connect(s, 0, avg, 1)
    g.connect(e, 0, avg, 2)
    neg = g.add_transform("negation")
    g.connect(avg, 0, neg, 0)
    first = g.compute(neg)
    calls = g.calls
    again = g.compute(neg)
    assert g.calls == calls and again.device_ptr == first.device_ptr  # cache hit: same handle, no copy, no launch
    launches = gpu.launch_count()
    g.compute(avg)
    assert gpu.launch_count() == launches
    g.set_constant(e, 48)
    third = g.compute(neg)
    assert g.calls == calls + 2  # average and negation re-ran; the cube constant did not re-upload
    gpu.set_sum_mode(gpu.SUM_SEQUENTIAL)
    try:
        g.set_constant(e, 40)
        assert_bit_equal(g.compute(neg).download(), -orc.average(host, 0, 40))
    finally:
        gpu.set_sum_mode(gpu.SUM_COMPENSATED)
    assert third.shape == (20, 24)
    # an error in an ancestor reaches the output with the node stack, and clears once the constant is fixed
    g.set_constant(e, 65)
    with pytest.raises(gpu.AflakError) as ei:
        g.compute(neg)
    assert "frame count (expected 65 < 65)" in ei.value.message and "average" in ei.value.message and "negation" in ei.value.message
    g.set_constant(e, 64)
    assert g.compute(neg).shape == (20, 24)


def test_concurrent_callers_on_their_own_streams(gpu, orc):
    """The ABI is entered from several rayon workers at once in aflak (dst/compute.rs:206-224): three threads reduce the
    same cached cube concurrently, each on its own stream, and a fourth thread consumes their outputs."""
    import threading

    host = orc.gen_cube(36, 96, 50, 52)
    cube = gpu.Image.upload(host)
    gpu.synchronize()
    out, errs = {}, []

    def work(i, s, e):
        try:
            gpu.set_device(0)
            gpu.set_sum_mode(gpu.SUM_SEQUENTIAL)
            for _ in range(10):
                out[i] = gpu.call("average", cube, s, e)[0]
        except Exception as ex:  # pragma: no cover
            errs.append(ex)

    ts = [threading.Thread(target=work, args=(i, s, e)) for i, (s, e) in enumerate([(0, 30), (30, 60), (60, 96)])]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    assert not errs, errs
    lc = gpu.call("linear_composition", out[0], out[1], 0.5, 0.5)  # consumer on the main thread's stream: event-ordered
    ew = gpu.call("create_equivalent_width", lc, out[2], 20.0, 1e10, True).download()
    off = orc.linear_composition(orc.average(host, 0, 30), orc.average(host, 30, 60), 0.5, 0.5)
    assert_bit_equal(ew, orc.equivalent_width(off, orc.average(host, 60, 96), 20.0, 1e10, True))


def test_fused_gather_single_rank(gpu, orc):
    """world = 1 degenerate case of the fused gather (aflak_b200/csrc/shard.cpp): band == whole image, flag wait passes."""
    nz, ny, nx = 100, 33, 40
    host = orc.gen_cube(37, nz, ny, nx)
    cube = gpu.Image.upload(host)
    gpu.synchronize()
    ga = gpu.Gather(1, 0, ny, nx)
    assert ga.band() == (0, ny)
    for it, (s, e) in enumerate([(0, nz), (10, 60), (99, 100)]):
        ga.reduce(cube.device_ptr, nz, [(s, e)], gpu.EPI_SUM)
        ga.wait()
        img = ga.result_image()
        s64, a64 = orc.sum_f64(host, s, e)
        assert_reduction(img.download(), s64, a64, f"gather step {it}")
    ga.check()
    ga.reduce(cube.device_ptr, nz, [(0, 50), (50, 100)], gpu.EPI_MEAN_RATIO, sum_mode=gpu.SUM_SEQUENTIAL)
    ga.wait()
    assert_bit_equal(ga.result_image().download(), orc.ratio(orc.average(host, 0, 50), orc.average(host, 50, 100)))
    # the two-slot protocol is enforced, not left to the caller: a reduce needs a wait before the next reduce
    ga.reduce(cube.device_ptr, nz, [(0, nz)], gpu.EPI_SUM)
    with pytest.raises(gpu.AflakError, match="afk_gather_wait has not been called"):
        ga.reduce(cube.device_ptr, nz, [(0, nz)], gpu.EPI_SUM)
    ga.wait()
    # EW needs its five parameters; the other epilogues read none
    with pytest.raises(ValueError):
        ga.reduce(cube.device_ptr, nz, [(0, 10), (10, 20), (20, 30)], gpu.EPI_EW, [1.0, 0.0, 20.0, 1e10])
    # a rejected call leaves the protocol where it was: the epoch only moves once the arguments have passed, so the next
    # reduce is accepted (no "wait has not been called") and its wait finds its own flag (no time-out, no NaN band)
    with pytest.raises(gpu.AflakError, match="Expected 0 <= start < end"):
        ga.reduce(cube.device_ptr, nz, [(10, nz + 1)], gpu.EPI_SUM)
    with pytest.raises(gpu.AflakError, match="Expected 0 <= start < end"):
        ga.reduce(cube.device_ptr, nz, [(7, 7)], gpu.EPI_SUM)
    ga.reduce(cube.device_ptr, nz, [(5, 25)], gpu.EPI_SUM, sum_mode=gpu.SUM_SEQUENTIAL)
    ga.wait(timeout_s=2.0)
    ga.check()
    assert_bit_equal(ga.result_image().download(), orc.integral(host, 5, 25), "reduce after two rejected calls")
    ga.reduce(cube.device_ptr, nz, [(0, nz)], gpu.EPI_SUM)
    ga.wait()
    # a consumer on another thread's stream: recorded on the handle, the next reduce is ordered after it
    import threading

    img = ga.result_image()
    box = {}
    t = threading.Thread(target=lambda: box.update(mm=gpu.call("image_min_max", img)))
    t.start()
    t.join()
    want = orc.integral(host, 0, nz) if gpu.get_sum_mode() == gpu.SUM_SEQUENTIAL else None
    s64, _ = orc.sum_f64(host, 0, nz)
    assert abs(float(box["mm"][0]) - s64.min()) <= 1e-5 * abs(s64.min()) and want is None
    ga.reduce(cube.device_ptr, nz, [(0, nz)], gpu.EPI_SUM)
    ga.wait()
    ga.check()
    ga.close()


def test_plain_c_client_of_the_abi(gpu, tmp_path):
    """The whole boundary from C, the way the Rust binding would drive it (tests/c/abi_client.c): registry walk, `integral` through
    the descriptor's function pointer with a closed-form answer, the reference's error wording, the evaluator's cache rule."""
    import json
    import subprocess

    from util import build_c_client

    p = subprocess.run([build_c_client(tmp_path)], capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stderr + p.stdout
    assert "abi_client ok" in p.stdout
    stats = json.loads(p.stdout.splitlines()[0])["c_abi_client"]
    assert stats["kernel_launches"] > 0
