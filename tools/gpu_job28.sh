#!/bin/bash
cat > /tmp/medt.py <<'PY'
import sys, os
sys.path.insert(0, os.getcwd())
import torch, aflak_b200 as afk
afk.set_device(0)
stream = torch.cuda.Stream(); torch.cuda.set_stream(stream); afk.set_stream(stream.cuda_stream)
for (nz, ny, nx) in [(3840, 1024, 1024), (2048, 2048, 2048), (8000, 512, 512)]:
    cube = afk.Image.synthetic(0xAF1A0003, nz, ny, nx)
    for _ in range(2): afk.call("median", cube, 0, nz)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(5): afk.call("median", cube, 0, nz)
    e1.record(stream); e1.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print(f"median {nz}x{ny}x{nx}: {ms:.3f} ms  {nz*ny*nx*4/ms/1e6:.0f} GB/s", flush=True)
    del cube
PY
timeout 300 python /tmp/medt.py 2>&1 | grep "median "
timeout 1200 python -m pytest tests/test_gpu_axis0.py -m gpu -q -x --timeout 900 -k median 2>&1 | tail -3
