#!/usr/bin/env python
"""Phase timing of the e2e step (pinned host cube -> upload -> fused reduce -> download)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import aflak_b200 as afk

afk.set_device(0)
nz, ny, nx = 3840, 1024, 1024
host_t = torch.empty((nz, ny, nx), dtype=torch.float32, pin_memory=True)
tmp = afk.Image.synthetic(3, nz, ny, nx)
afk.lib.afk_image_download(tmp.handle, host_t.data_ptr())
del tmp
host = host_t.numpy()
res = np.empty((ny, nx), np.float32)
for it in range(5):
    afk.synchronize()
    t0 = time.perf_counter()
    img = afk.Image.upload(host, sync=False)
    t1 = time.perf_counter()
    afk.synchronize()
    t2 = time.perf_counter()
    out = afk.fused_average_ratio(img, 0, nz // 2, nz // 2, nz)
    afk.synchronize()
    t3 = time.perf_counter()
    out.download(out=res)
    t4 = time.perf_counter()
    del img, out
    afk.synchronize()
    t5 = time.perf_counter()
    print(f"iter {it}: upload-call {1e3*(t1-t0):.1f} ms, upload-done {1e3*(t2-t0):.1f} ms ({host.nbytes/(t2-t0)/1e9:.1f} GB/s), "
          f"reduce {1e3*(t3-t2):.2f} ms, download {1e3*(t4-t3):.2f} ms, free {1e3*(t5-t4):.2f} ms", flush=True)
# torch reference copy of the same pinned buffer
d = torch.empty((nz, ny, nx), dtype=torch.float32, device="cuda")
for it in range(3):
    torch.cuda.synchronize(); t0 = time.perf_counter(); d.copy_(host_t, non_blocking=True); torch.cuda.synchronize()
    print(f"torch copy_: {host.nbytes/(time.perf_counter()-t0)/1e9:.1f} GB/s")
