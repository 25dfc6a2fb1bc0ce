"""CPU: the C-ABI library loads, exports every symbol include/aflak_b200.h declares, and its host-side logic
(registry, value model, conversions, colour-LUT tables, unit algebra, graph bookkeeping, band partition) behaves like
the reference interface it replaces.  No compute call succeeds without a GPU -- and that failure is loud.

Reference: aflak_primitives/src/lib.rs:66-136,141-823; aflak_cake/src/transform.rs, tests/{algo,dst}.rs;
aflak_plot/src/imshow/lut.rs:292-343.
"""
import ctypes as C
import json
import os
import re

import numpy as np
import pytest

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_every_declared_symbol_is_exported(afk):
    src = open(afk.HEADER_PATH).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    names = set(re.findall(r"\b(afk_[a-z0-9_]+)\s*\(", src))
    names -= {"afk_plain_function", "afk_status"}  # the function-pointer typedef
    assert len(names) > 90
    missing = [n for n in sorted(names) if not hasattr(afk.lib, n)]
    assert not missing, missing
    assert afk.lib.afk_abi_version() == 1


def test_no_oracle_in_product(afk):
    """The product must not link, import or call the oracle (it is the checker)."""
    root = os.path.dirname(afk.HEADER_PATH)
    pkg = os.path.join(os.path.dirname(root), "aflak_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cpp", ".cu", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in text and "aflak_oracle" not in text.replace("oracle/aflak_oracle.c orc_gen_voxel", ""), f


# name -> (version, [(slot, type, default)], [outputs])   from aflak_primitives/src/lib.rs:141-747
REGISTRY = {
    "integral": ((1, 0, 0), [("image", "Image", None), ("start", "Integer", 0), ("end", "Integer", 1)], ["Image", "Float", "Float"]),
    "average": ((1, 0, 0), [("image", "Image", None), ("start", "Integer", 0), ("end", "Integer", 1)], ["Image", "Float", "Float"]),
    "slice_one_frame": ((1, 0, 0), [("image", "Image", None), ("frame", "Integer", 0)], ["Image"]),
    "linear_composition": ((1, 0, 0), [("u", "Image", None), ("v", "Image", None), ("a", "Float", 1.0), ("b", "Float", 1.0)], ["Image"]),
    "image_multiplier": ((1, 0, 0), [("u", "Image", None), ("v", "Image", None), ("a", "Float", 1.0), ("b", "Float", 1.0)], ["Image"]),
    "create_equivalent_width": ((0, 1, 0), [("i_off", "Image", None), ("i_on", "Image", None), ("fl", "Float", 1.0),
                                            ("max", "Float", float("inf")), ("is_emission", "Bool", False)], ["Image"]),
    "down_sampling": ((1, 0, 0), [("image", "Image", None), ("n", "Integer", 1)], ["Image"]),
    "convert_to_logscale": ((1, 0, 0), [("image", "Image", None), ("a", "Float", 1000.0), ("v_min", "Float", None), ("v_max", "Float", None)], ["Image"]),
    "log10": ((1, 0, 0), [("image", "Image", None)], ["Image"]),
    "apply_tone_curve": ((1, 0, 0), [("image", "Image", None), ("tone_curve", "ToneCurve", None)], ["Image"]),
    "image_min_max": ((1, 0, 0), [("image", "Image", None)], ["Float", "Float"]),
    "open_fits": ((1, 0, 0), [("path", "Paths", None), ("n", "Integer", 0)], ["Fits"]),
    "fits_to_image": ((1, 0, 0), [("fits", "Fits", None), ("hdu_idx", "Integer", 0), ("extension", "Str", "")], ["Image"]),
    "make_float3": ((1, 0, 0), [("f1", "Float", 0.0), ("f2", "Float", 0.0), ("f3", "Float", 0.0)], ["Float3"]),
    "make_plane3d": ((1, 0, 0), [("p0", "Float3", (0.0, 0.0, 0.0)), ("dir1", "Float3", (0.0, 0.0, 1.0)), ("dir2", "Float3", (0.0, 1.0, 0.0)), ("count1", "Integer", 1),
                                 ("count2", "Integer", 1)], ["Map2dTo3dCoords"]),
    "slice_3d_to_2d": ((0, 1, 0), [("image", "Image", None), ("map", "Map2dTo3dCoords", None)], ["Image"]),
    "extrude": ((0, 1, 0), [("image", "Image", None), ("roi", "Roi", "All")], ["Image"]),
    "extract_wave": ((1, 0, 0), [("image", "Image", None), ("roi", "Roi", "All")], ["Image"]),
    "create_velocity_field_map": ((1, 0, 0), [("image", "Image", None), ("w_0", "Float", 0.0)], ["Image"]),
    "color_image_to_hsv": ((1, 0, 0), [("image", "Image", None)], ["Image"]),
    "color_image_from_rgb": ((1, 0, 0), [("image_r", "Image", None), ("image_g", "Image", None), ("image_b", "Image", None)], ["Image"]),
    "histogram_transformation": ((0, 1, 0), [("image", "Image", None), ("parameters", "Float3", (0.0, 0.5, 1.0)), ("_dynamic_range", "Float2", (0.0, 1.0))], ["Image"]),
    "histogram_transformation_rgb": ((0, 1, 0), [("image", "Image", None), ("parameters", "Float3x3", (0.0, 0.5, 1.0) * 3), ("_dynamic_range", "Float2", (0.0, 1.0))], ["Image"]),
    "apply_arcsinh_stretch": ((0, 1, 0), [("image", "Image", None), ("beta", "Float", float(np.float32(0.000000001)))], ["Image"]),
    "variance": ((1, 0, 0), [("image", "Image", None), ("start", "Integer", 0), ("end", "Integer", 1)], ["Image"]),
    "stddev": ((1, 0, 0), [("image", "Image", None), ("start", "Integer", 0), ("end", "Integer", 1)], ["Image"]),
    "median": ((1, 0, 0), [("image", "Image", None), ("start", "Integer", 0), ("end", "Integer", 1)], ["Image"]),
    "extract_argmin_max_wavelength": ((0, 1, 0), [("image", "Image", None), ("start", "Integer", 0), ("end", "Integer", 1),
                                                  ("is_min", "Bool", False)], ["Image", "Image"]),
    "peak_based_wavelength_range": ((0, 1, 0), [("image", "Image", None), ("start", "Integer", 0), ("end", "Integer", 1),
                                                ("range", "Integer", 1), ("is_min", "Bool", False)], ["Image", "Image"]),
    "extract_centrobaric_wavelength": ((0, 1, 0), [("image", "Image", None), ("start", "Integer", 0), ("end", "Integer", 1)], ["Image"]),
    "extract_centrobaric_wavelength_with_mask": ((0, 1, 0), [("image", "Image", None), ("start_mask", "Image", None),
                                                             ("end_mask", "Image", None)], ["Image"]),
    "ratio_from_bands": ((1, 0, 0), [("z", "Float", None), ("z1", "Float", None), ("z2", "Float", None)], ["Float", "Float"]),
    "range_specification": ((0, 1, 0), [("image", "Image", None), ("start", "Integer", 0), ("end", "Integer", 1)], ["Image"]),
}


def test_registry_matches_reference(afk):
    names = [t.name for t in afk.transformations()]
    assert len(names) == len(set(names))
    for name, (ver, slots, outs) in REGISTRY.items():
        t = afk.get_transform(name)
        assert t is not None, name
        assert t.version == ver, name
        assert t.outputs == outs, name
        assert [(s[0], s[1]) for s in t.inputs] == [(s[0], s[1]) for s in slots], name
        for got, want in zip(t.inputs, slots):
            if want[2] is None:
                assert got[2] is None or want[1] == "ToneCurve", (name, got)
            elif want[1] == "Roi":
                assert got[2] == afk.Roi(), (name, got)  # roi::ROI::All
            elif isinstance(want[2], tuple):
                assert np.array_equal(got[2], np.array(want[2], dtype=np.float32)), (name, got, want)
            else:
                assert got[2] == want[2], (name, got, want)
    assert afk.get_transform("no_such_transform") is None


def test_type_ids_and_conversions(afk):
    want = ["Integer", "Float", "Float2", "Float3", "Float3x3", "Str", "Bool", "Fits", "Image", "Map2dTo3dCoords", "Roi", "Paths", "ToneCurve"]
    assert [afk.type_name(i) for i in range(13)] == want
    for i, n in enumerate(want):
        assert afk.lib.afk_type_from_name(n.encode()) == i
    assert afk.lib.afk_type_from_name(b"Nope") == -1
    # ConvertibleVariants table (lib.rs:795-823): Integer <-> Float only
    for a in range(13):
        for b in range(13):
            assert bool(afk.lib.afk_type_convertible(a, b)) == (a == b or {a, b} == {afk.T_INTEGER, afk.T_FLOAT})
    assert not afk.lib.afk_type_editable(afk.T_IMAGE) and not afk.lib.afk_type_editable(afk.T_FITS)
    assert afk.lib.afk_type_editable(afk.T_TONECURVE)


def _convert(afk, x, into):
    v = afk._to_value(x)
    try:
        c = afk.lib.afk_value_convert(C.c_void_p(v), into)
        if not c:
            return None
        try:
            return afk._from_value(C.c_void_p(c))
        finally:
            afk.lib.afk_value_release(C.c_void_p(c))
    finally:
        afk.lib.afk_value_release(C.c_void_p(v))


def test_value_model(afk):
    # aflak_cake/tests/algo.rs:25: Integer -> Float auto-conversion; lib.rs:817-823 Float -> Integer is round()
    assert _convert(afk, 1, afk.T_FLOAT) == np.float32(1.0)
    assert _convert(afk, 2.5, afk.T_INTEGER) == 3 and _convert(afk, -2.5, afk.T_INTEGER) == -3
    assert _convert(afk, 2.4, afk.T_INTEGER) == 2
    assert _convert(afk, float("nan"), afk.T_INTEGER) == 0
    assert _convert(afk, 1e30, afk.T_INTEGER) == 2 ** 63 - 1
    assert _convert(afk, "s", afk.T_FLOAT) is None
    # defaults (lib.rs:760-776)
    for t, want in [(afk.T_INTEGER, 0), (afk.T_FLOAT, 0.0), (afk.T_STR, ""), (afk.T_BOOL, False)]:
        v = afk.lib.afk_value_default_for(t)
        assert afk._from_value(C.c_void_p(v)) == want
        afk.lib.afk_value_release(C.c_void_p(v))
    v = afk.lib.afk_value_default_for(afk.T_TONECURVE)
    tc = afk._from_value(C.c_void_p(v))
    assert tc.arr.size == 256 and np.array_equal(tc.arr, np.arange(256, dtype=np.float32))
    afk.lib.afk_value_release(C.c_void_p(v))
    assert not afk.lib.afk_value_default_for(afk.T_IMAGE)
    # PartialEq
    a, b, c = afk._to_value(3), afk._to_value(3), afk._to_value(3.0)
    assert afk.lib.afk_value_eq(C.c_void_p(a), C.c_void_p(b)) == 1
    assert afk.lib.afk_value_eq(C.c_void_p(a), C.c_void_p(c)) == 0
    for v in (a, b, c):
        afk.lib.afk_value_release(C.c_void_p(v))
    f3 = afk._to_value((1.0, 2.0, 3.0))
    assert afk.lib.afk_value_type(C.c_void_p(f3)) == afk.T_FLOAT3
    assert afk._from_value(C.c_void_p(f3)).tolist() == [1.0, 2.0, 3.0]
    afk.lib.afk_value_release(C.c_void_p(f3))


def test_host_only_transform_and_errors(afk):
    """ratio_from_bands is pure host arithmetic (lib.rs:226-246): callable without a GPU, through feed + call."""
    k = json.load(open(os.path.join(GOLD, "kat.json")))
    r = k["ratio_from_bands"]
    r2, r1 = afk.call("ratio_from_bands", r["z"], r["z1"], r["z2"])
    assert (r2, r1) == (np.float32(r["off_ratio_2"]), np.float32(r["off_ratio_1"]))
    # Integer arguments are converted to Float by feed (transform.rs:330-346)
    assert afk.call("ratio_from_bands", 3144, 3109, 3184) == (r2, r1)
    res = afk.get_transform("ratio_from_bands").call_results(3.0, 5.0, 9.0)
    assert all(isinstance(x, afk.AflakError) and x.status == afk.ERR_UNEXPECTED_INPUT for x in res)
    assert res[0].message == k["errors"]["ratio_from_bands_order"]
    # wrong argument type: cake reports a conversion error before calling f
    res = afk.get_transform("ratio_from_bands").call_results("x", 5.0, 9.0)
    assert res[0].status == afk.ERR_ARGUMENT and "Cannot convert 'Str' to 'Float'" in res[0].message
    # missing argument without default
    res = afk.get_transform("ratio_from_bands").call_results(1.0, None, 9.0)
    assert res[0].status == afk.ERR_ARGUMENT


def test_lut_host_side_golden(afk):
    """ColorLUT table construction and color_at are host code in the product too: pin them on the reference's vectors."""
    gold = json.load(open(os.path.join(GOLD, "lut_golden.json")))
    for case in gold["color_at"]:
        lut = afk.Lut([(s[0], tuple(s[1])) for s in case["gradient"]])
        if "table_lims" in case:  # lims assigned after the table was built (lut.rs:332-334)
            lut.set_lims(*case["lims"])
        for point, rgb in case["cases"]:
            assert lut.color_at(point) == rgb, (point, rgb)
    b = gold["bounds"]
    lut = afk.Lut([(s[0], tuple(s[1])) for s in b["gradient"]])
    want = [((p[0][0], tuple(p[0][1])), (p[1][0], tuple(p[1][1]))) for p in b["pairs"]]
    assert lut.bounds() == want


def test_lut_tables_match_oracle(afk, orc):
    assert afk.Lut.builtin_names() == ["Grey", "GreyClip", "Thermal", "Flame", "Yellowy", "HeatMap", "HeatMap_Inv", "Red", "Blue", "Green"]
    for name in afk.Lut.builtin_names():
        assert np.array_equal(afk.Lut(name).table(), orc.lut_build(name)), name
    lut = afk.Lut("Flame")
    table = orc.lut_build("Flame")
    lut.set_lims(0.2, 0.3, 0.9)
    rng = np.random.default_rng(0)
    for p in np.concatenate([rng.uniform(-0.2, 1.2, 200), [0.0, 0.2, 0.9, 1.0, np.nan]]).astype(np.float32):
        assert lut.color_at(float(p)) == orc.color_at(table, float(p), (0.2, 0.3, 0.9)), p
        assert lut.color_at_bounds(float(p) * 10 - 3, -3.0, 7.0) == orc.color_at_bounds(table, float(p) * 10 - 3, -3.0, 7.0, (0.2, 0.3, 0.9))
    with pytest.raises(afk.AflakError):
        afk.Lut("NoSuchGradient")


def test_shard_bands_tile_the_image(afk):
    for ny in (0, 1, 7, 74, 1024, 4096, 4099):
        for world in (1, 2, 3, 4, 8):
            bands = [afk.shard_band(ny, world, r) for r in range(world)]
            assert bands[0][0] == 0 and sum(b[1] for b in bands) == ny
            for (y0, rows), (y1, _) in zip(bands, bands[1:]):
                assert y0 + rows == y1
            assert max(b[1] for b in bands) - min(b[1] for b in bands) <= 1
    with pytest.raises(afk.AflakError):
        afk.shard_band(10, 2, 2)


def test_graph_bookkeeping_without_gpu(afk):
    """Port of aflak_cake/tests/dst.rs on host-only nodes: type check, cycle check, memoisation, invalidation."""
    g = afk.Graph()
    z, z1, z2 = g.add_constant(3144.0), g.add_constant(3109.0), g.add_constant(3184.0)
    rb = g.add_transform("ratio_from_bands")
    for i, c in enumerate((z, z1, z2)):
        g.connect(c, 0, rb, i)
    r2 = g.compute(rb, 0)
    r1 = g.compute(rb, 1)
    assert np.float32(r1) == np.float32(np.float32(35.0) / np.float32(75.0)) and r2 == np.float32(1.0) - r1
    calls = g.calls
    g.compute(rb, 0)
    assert g.calls == calls  # cache hit (cache.rs:130-136)
    g.set_constant(z, 3150.0)  # update_constant_node bumps updated_on -> recompute (dst.rs:114-143)
    assert np.float32(g.compute(rb, 1)) == np.float32(np.float32(41.0) / np.float32(75.0))
    assert g.calls == calls + 1
    # IncompatibleTypes (dst.rs:96): Str constant into a Float slot
    s = g.add_constant("text")
    with pytest.raises(afk.AflakError) as ei:
        g.connect(s, 0, rb, 0)
    assert "does not provide the required input type" in ei.value.message
    # convertible edge: Integer constant into a Float slot (dst.rs:96-112)
    i = g.add_constant(3160)
    g.connect(i, 0, rb, 0)
    assert np.float32(g.compute(rb, 1)) == np.float32(np.float32(51.0) / np.float32(75.0))
    # cycle: feed rb's own output into its input
    rb2 = g.add_transform("ratio_from_bands")
    g.connect(rb, 0, rb2, 0)
    with pytest.raises(afk.AflakError) as ei:
        g.connect(rb2, 0, rb, 1)
    assert "cycle" in ei.value.message
    # missing dependency
    with pytest.raises(afk.AflakError) as ei:
        g.compute(rb2, 0)
    assert ei.value.status == afk.ERR_ARGUMENT
    # errors propagate down the graph with the node stack (dst/compute.rs:240-247)
    g.disconnect(rb, 0)
    bad = g.add_constant(1.0)
    g.connect(bad, 0, rb, 0)
    with pytest.raises(afk.AflakError) as ei:
        g.compute(rb, 0)
    assert "wrong magnitude correlation" in ei.value.message and "ratio_from_bands" in ei.value.message


def test_compute_entry_points_fail_loudly_without_gpu(afk):
    if afk.device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(afk.AflakError) as ei:
        afk.Image.upload(np.zeros((2, 2), np.float32))
    assert ei.value.status == afk.ERR_CUDA and "no CPU fallback" in ei.value.message
    with pytest.raises(afk.AflakError):
        afk.Image.synthetic(1, 4, 4, 4)
    with pytest.raises(afk.AflakError):
        afk.set_device(0)


def test_wrap_device_validates_its_shape_before_anything_else(afk):
    """afk_image_wrap_device takes extents from the caller: negative or overflowing ones are refused (the kernels index with
    their product) -- checked before the device is touched, so this runs without a GPU."""
    import ctypes as C

    fake = C.c_void_p(0x1000)  # never dereferenced
    for shape in ([4, -1], [1 << 40, 1 << 40], [-5]):
        sh = (C.c_int64 * len(shape))(*shape)
        assert not afk.lib.afk_image_wrap_device(fake, len(shape), sh), shape
        msg = afk.lib.afk_last_error().decode()
        assert "negative dimension" in msg or "image too large" in msg, msg
    assert not afk.lib.afk_image_wrap_device(fake, 2, None)
    assert "bad arguments" in afk.lib.afk_last_error().decode()
    assert not afk.lib.afk_image_wrap_device(None, 0, None)


def test_header_is_plain_c_and_a_c_client_links(afk, tmp_path):
    """include/aflak_b200.h must be consumable by a C compiler (it is what a Rust `-sys` crate binds): the C99 client of
    tests/c/ compiles with -Wall -Wextra -Werror, links against the shared library and, without a GPU, stops at its device check."""
    import subprocess

    from util import build_c_client

    exe = build_c_client(tmp_path)
    p = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    import aflak_b200

    if aflak_b200.device_count() == 0:
        assert p.returncode == 77 and "no CUDA device" in p.stderr
    else:
        assert p.returncode == 0, p.stderr


def test_sum_mode_is_a_process_default_with_a_thread_override(afk):
    """cake calls operators from arbitrary rayon workers (aflak_cake/src/dst/compute.rs:206-224): the host sets the mode
    once for the process; a thread may override it for itself."""
    import threading

    seen = {}

    def worker(name, override=None):
        if override is not None:
            afk.set_sum_mode(override)
        seen[name] = afk.get_sum_mode()

    assert afk.get_sum_mode() == afk.SUM_COMPENSATED
    try:
        afk.set_default_sum_mode(afk.SUM_SEQUENTIAL)
        assert afk.get_sum_mode() == afk.SUM_SEQUENTIAL  # no override on this thread
        for name, ov in (("plain", None), ("override", afk.SUM_COMPENSATED)):
            t = threading.Thread(target=worker, args=(name, ov))
            t.start()
            t.join()
        assert seen == {"plain": afk.SUM_SEQUENTIAL, "override": afk.SUM_COMPENSATED}
        afk.set_sum_mode(afk.SUM_COMPENSATED)
        assert afk.get_sum_mode() == afk.SUM_COMPENSATED
        afk.set_sum_mode(afk.SUM_INHERIT)
        assert afk.get_sum_mode() == afk.SUM_SEQUENTIAL
    finally:
        afk.set_default_sum_mode(afk.SUM_COMPENSATED)
        afk.set_sum_mode(afk.SUM_INHERIT)
    with pytest.raises(afk.AflakError):
        afk.set_default_sum_mode(7)


def test_tone_curve_from_points_host(afk, orc):
    """afk_tone_curve_from_points restates UiToneCurve::create_curve (imgui_tone_curve/src/lib.rs:68-158) in the product
    (host code, no device needed): closed forms for 2 and 3 points, the oracle's restatement for the spline cases, the
    reference's panic for a NaN abscissa."""
    arr = afk.tone_curve_from_points([(0.0, 0.0), (1.0, 1.0)])
    assert arr.dtype == np.float32 and np.array_equal(arr, np.arange(256, dtype=np.float32))  # ToneCurveState::default's table
    arr = afk.tone_curve_from_points([(0.0, 0.25), (0.5, 0.75)])  # slope 1, offset 0.25: (i/256 + 0.25) * 256
    assert np.array_equal(arr, np.arange(256, dtype=np.float32) + np.float32(64.0))
    # three points on y = x^2: a = 1, b = c = 0 up to rounding -> (t*t)*256 within a few ulps of the quadratic's own rounding
    arr = afk.tone_curve_from_points([(0.0, 0.0), (0.5, 0.25), (1.0, 1.0)])
    t = np.arange(256, dtype=np.float32) / np.float32(256.0)
    assert np.allclose(arr, t * t * 256.0, rtol=0, atol=1e-4)
    rng = np.random.default_rng(0xAF1A0090)
    cases = [[(1.0, 1.0), (0.0, 0.0), (0.3, 0.6)], [(0, 0), (0.25, 0.4), (0.6, 0.5), (1, 1)], [(0, 0), (0.2, 0.1), (0.5, 0.8), (0.7, 0.75), (1, 1)],
             [(0.0, 0.0)], [(0, 0), (0.5, 0.5), (0.5, 0.6), (1, 1)]]
    cases += [[(float(x), float(y)) for x, y in rng.random((n, 2))] for n in (3, 4, 5, 6, 9, 12)]
    for cp in cases:
        got = afk.tone_curve_from_points(cp)
        want = np.asarray(orc.create_curve(cp), dtype=np.float32)
        assert got.shape == want.shape, (cp, got.shape, want.shape)
        nan = np.isnan(want)
        assert np.array_equal(np.isnan(got), nan) and np.array_equal(got[~nan].view(np.uint32), want[~nan].view(np.uint32)), cp
    assert afk.tone_curve_from_points(np.zeros((0, 2), np.float32)).size == 0
    with pytest.raises(afk.AflakError) as ei:
        afk.tone_curve_from_points([(0.0, 0.0), (float("nan"), 0.5), (1.0, 1.0)])
    assert ei.value.status == afk.ERR_PANIC
