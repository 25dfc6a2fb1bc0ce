"""CPU, world_size 2 over gloo: the host-side logic of the sharded path (one process per GPU).

What runs here without a GPU: the band partition (afk_shard_band through the C ABI), the exchange of the 64-byte
handles through the process group (the only thing that ever crosses it), and the layout rule "rank r's result band is
rows [y0, y0+rows) of the full image".  The per-band reduction is done by the ORACLE here (test infrastructure standing
in for the kernel, which needs a GPU); the fused peer-store gather itself is covered on GPUs by bench.py --gpus N and
tests/test_gpu_graph.py::test_fused_gather_single_rank.
"""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, shape, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch
    import torch.distributed as dist

    import aflak_b200 as afk
    import oracle

    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        nz, ny, nx = shape
        y0, rows = afk.shard_band(ny, world, rank)
        # every rank regenerates only its own band of the global cube (counter-based generator keyed by global index)
        band = oracle.gen_cube(0xAF1A0005, nz, ny, nx, y0=y0, rows=rows, threads=1)
        a0 = oracle.average(band, 0, nz // 2)
        a1 = oracle.average(band, nz // 2, nz)
        part = oracle.ratio(a0, a1)
        # handle exchange, as bench.py does for afk_gather_connect: 64 opaque bytes per rank, rank-major
        mine = bytes([rank]) * 64
        handles = [None] * world
        dist.all_gather_object(handles, mine)
        assert handles == [bytes([r]) * 64 for r in range(world)]
        # gather of the 2-D result: bands may differ by one row -> pad to the widest band
        maxrows = max(afk.shard_band(ny, world, r)[1] for r in range(world))
        buf = torch.zeros((maxrows, nx), dtype=torch.float32)
        buf[:rows] = torch.from_numpy(part)
        out = [torch.zeros_like(buf) for _ in range(world)]
        dist.all_gather(out, buf)
        full = np.concatenate([out[r][: afk.shard_band(ny, world, r)[1]].numpy() for r in range(world)], axis=0)
        # global min/max of the result (image_min_max feeding convert_to_logscale): all-reduce of one float each
        mn, mx = oracle.image_min_max(part)
        t = torch.tensor([-float(mn), float(mx)])
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if rank == 0:
            whole = oracle.gen_cube(0xAF1A0005, nz, ny, nx, threads=1)
            want = oracle.ratio(oracle.average(whole, 0, nz // 2), oracle.average(whole, nz // 2, nz))
            wmn, wmx = oracle.image_min_max(want)
            q.put((bool(np.array_equal(full, want)), full.shape == (ny, nx), float(-t[0]) == float(wmn) and float(t[1]) == float(wmx)))
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("shape", [(24, 37, 20), (16, 8, 33)])
def test_sharded_average_ratio_world2(shape):
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, shape, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=180)
    for p in procs:
        assert p.exitcode == 0, p.exitcode
    same, shape_ok, minmax_ok = q.get(timeout=10)
    assert same and shape_ok and minmax_ok
