"""GPU: afk_k_reduce_static (the allocation-free variant of afk_k_reduce, aflak_b200/csrc/noalloc.cu) against afk_k_reduce.

No counterpart in the reference: the entry point was added for callers that cannot call
cudaMalloc (CUDA graph capture, tight serving loops).  Same kernels and launch plan as afk_k_reduce, so the bar is the one the
reduction tests use for identical arithmetic: bit for bit (tests/util.py assert_bit_equal), on two calls in a row -- the static
scratch keeps the first call's contents.  This file is named to be collected last.
"""
import numpy as np
import pytest

from util import assert_bit_equal

pytestmark = pytest.mark.gpu

# (nz, ny, nx, the three-window plan of the default sum mode fits the static scratch, why)
SHAPES = [
    (200, 512, 512, False, "256 pixel tiles shared between 444 CTAs: split-K partials and arrival counters in the static scratch"),
    (457, 74, 74, True, "MaNGA spatial size: small-image variants (k-groups / thread-block cluster)"),
    (64, 128, 128, True, "small tiles, three windows fit the scratch"),
    (33, 37, 53, True, "odd pixel count: scalar lanes"),
]


def _cases(gpu, nz):
    a, b, c = nz // 5, nz // 2, nz - 3
    yield "integral", gpu.EPI_SUM, [(0, nz)], None
    yield "average", gpu.EPI_MEAN, [(a, c)], None
    yield "average/average ratio", gpu.EPI_MEAN_RATIO, [(0, b), (b, nz)], None
    yield "equivalent width", gpu.EPI_EW, [(b - 4, b + 4), (a, a + 6), (c - 6, c)], [0.5, 0.5, 20.0, 1e10, -1.0]


def _run(fn, gpu, cube, nz, n2, shape2d, windows, epi, params, mode, **kw):
    out = gpu.Image.alloc(shape2d)
    fn(cube.device_ptr, nz, n2, windows, epi, out.device_ptr, params=params, sum_mode=mode, **kw)
    return out


def test_k_reduce_static_matches_k_reduce_twice_in_a_row(gpu):
    import torch

    stream = torch.cuda.Stream()
    for nz, ny, nx, ew_fits, why in SHAPES:
        cube = gpu.Image.synthetic(0xAF1A0051, nz, ny, nx, pathology=(ny == 74))
        n2 = ny * nx
        for mode in (gpu.SUM_COMPENSATED, gpu.SUM_SEQUENTIAL):
            for name, epi, windows, params in _cases(gpu, nz):
                want = _run(gpu.k_reduce, gpu, cube, nz, n2, (ny, nx), windows, epi, params, mode)
                # compensated mode splits 256 tiles over 427 CTAs: 427 x 2 slots x 3 windows x 16 KiB = 42 MB of partials > 32 MiB;
                # the sequential plan (whole tiles only, 256 CTAs, 25 MB) fits
                if epi == gpu.EPI_EW and not ew_fits and mode == gpu.SUM_COMPENSATED:
                    with pytest.raises(gpu.AflakError) as e:
                        gpu.k_reduce_static(cube.device_ptr, nz, n2, windows, epi, want.device_ptr, params=params, sum_mode=mode)
                    assert e.value.status == gpu.ERR_UNEXPECTED_INPUT and "static scratch" in str(e.value), str(e.value)
                    continue
                first, second = gpu.Image.alloc((ny, nx)), gpu.Image.alloc((ny, nx))
                gpu.synchronize()  # cube, `want` and the (stream-ordered) output allocations are ready; the static calls run on another stream
                for out in (first, second):
                    gpu.k_reduce_static(cube.device_ptr, nz, n2, windows, epi, out.device_ptr, params=params, sum_mode=mode, stream=stream.cuda_stream)
                stream.synchronize()
                w = want.download()
                assert_bit_equal(first.download(), w, f"{name} {nz}x{ny}x{nx} mode {mode}, first call ({why})")
                assert_bit_equal(second.download(), w, f"{name} {nz}x{ny}x{nx} mode {mode}, second call ({why})")
        del cube


def test_k_reduce_static_allocates_nothing_and_survives_stale_scratch(gpu):
    """Free device memory does not move across the calls, and a call with another plan in between (which leaves its partials in
    the scratch) does not change the answer."""
    import torch

    nz, ny, nx = 120, 512, 512
    cube = gpu.Image.synthetic(0xAF1A0052, nz, ny, nx)
    n2 = ny * nx
    win = [(0, nz // 2), (nz // 2, nz)]
    want = _run(gpu.k_reduce, gpu, cube, nz, n2, (ny, nx), win, gpu.EPI_MEAN_RATIO, None, gpu.SUM_COMPENSATED).download()
    outs = [gpu.Image.alloc((ny, nx)) for _ in range(3)]
    gpu.synchronize()
    gpu.k_reduce_static(cube.device_ptr, nz, n2, win, gpu.EPI_MEAN_RATIO, outs[0].device_ptr)  # set-up call (resolves the array)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    launches0 = gpu.launch_count()
    gpu.k_reduce_static(cube.device_ptr, nz, n2, win, gpu.EPI_MEAN_RATIO, outs[1].device_ptr)
    gpu.k_reduce_static(cube.device_ptr, nz, n2, [(3, nz - 1)], gpu.EPI_SUM, outs[2].device_ptr)  # another plan over the same scratch
    gpu.k_reduce_static(cube.device_ptr, nz, n2, win, gpu.EPI_MEAN_RATIO, outs[2].device_ptr)
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free0, "afk_k_reduce_static changed the amount of free device memory"
    assert gpu.launch_count() - launches0 == 3, "one kernel launch per call"
    for i, o in enumerate(outs):
        assert_bit_equal(o.download(), want, f"call {i}")


def test_k_reduce_static_rejects_plans_beyond_the_scratch(gpu):
    """Three windows over a megapixel image: 444 CTAs x 2 slots x 3 windows x 16 KiB of partials = 43.6 MB > 32 MiB."""
    nz, ny, nx = 40, 1024, 1024
    cube = gpu.Image.synthetic(0xAF1A0053, nz, ny, nx)
    out = gpu.Image.alloc((ny, nx))
    with pytest.raises(gpu.AflakError) as e:
        gpu.k_reduce_static(cube.device_ptr, nz, ny * nx, [(10, 20), (0, 8), (30, 38)], gpu.EPI_EW, out.device_ptr, params=[0.5, 0.5, 20.0, 1e10, -1.0])
    assert e.value.status == gpu.ERR_UNEXPECTED_INPUT
    assert "static scratch" in str(e.value) and "33554432" in str(e.value)
    # the two-window case of the same image fits (29.1 MB) and matches
    win = [(0, 20), (20, 40)]
    want = gpu.Image.alloc((ny, nx))
    gpu.k_reduce(cube.device_ptr, nz, ny * nx, win, gpu.EPI_MEAN_RATIO, want.device_ptr)
    gpu.synchronize()
    gpu.k_reduce_static(cube.device_ptr, nz, ny * nx, win, gpu.EPI_MEAN_RATIO, out.device_ptr)
    import torch

    torch.cuda.synchronize()
    assert_bit_equal(out.download(), want.download(), "1024x1024 two windows")
    # argument checks keep afk_k_reduce's wording
    with pytest.raises(gpu.AflakError) as e:
        gpu.k_reduce_static(cube.device_ptr, nz, ny * nx, [(5, 5)], gpu.EPI_SUM, out.device_ptr)
    assert "Expected start < end" in str(e.value)
