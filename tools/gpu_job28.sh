#!/bin/bash
cat > /tmp/medt.py <<'PY'
import sys, os
sys.path.insert(0, os.getcwd())
import torch, aflak_b200 as afk
afk.set_device(0)
stream = torch.cuda.Stream(); torch.cuda.set_stream(stream); afk.set_stream(stream.cuda_stream)
for (nz, ny, nx) in [(40, 2048, 2048), (65, 2048, 2048), (100, 2048, 2048), (128, 2048, 2048), (129, 2048, 2048), (200, 2048, 2048)]:
    cube = afk.Image.synthetic(0xAF1A0003, nz, ny, nx)
    for _ in range(2): afk.call("median", cube, 0, nz)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(5): afk.call("median", cube, 0, nz)
    e1.record(stream); e1.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print(f"median min={os.environ.get('AFK_MEDIAN_STREAM_MIN','dflt')} {nz}x{ny}x{nx}: {ms:.3f} ms  {nz*ny*nx*4/ms/1e6:.0f} GB/s", flush=True)
    del cube
PY
for m in 65; do AFK_MEDIAN_STREAM_MIN=$m timeout 300 python /tmp/medt.py 2>&1 | grep "median "; done
