#!/usr/bin/env python
"""Phase timing of the e2e step (pinned host cube -> upload -> fused reduce -> download), with variations."""
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


def run(label, keep_old, torch_stream):
    if torch_stream:
        st = torch.cuda.Stream()
        torch.cuda.set_stream(st)
        afk.set_stream(st.cuda_stream)
    else:
        afk.set_stream(0)
    old = None
    for it in range(5):
        afk.synchronize()
        t0 = time.perf_counter()
        img = afk.Image.upload(host, sync=False)
        t1 = time.perf_counter()
        afk.synchronize()
        t2 = time.perf_counter()
        out = afk.fused_average_ratio(img, 0, nz // 2, nz // 2, nz)
        out.download(out=res)
        t3 = time.perf_counter()
        if keep_old:
            old = img  # the previous cube stays alive until after the next upload (as in a graph constant)
        else:
            del img
        del out
        print(f"{label} iter {it}: upload-call {1e3*(t1-t0):.1f} ms, upload-done {1e3*(t2-t0):.1f} ms, reduce+download {1e3*(t3-t2):.2f} ms", flush=True)
    del old


run("own-stream/free-first", False, False)
run("own-stream/keep-old", True, False)
run("torch-stream/free-first", False, True)
run("torch-stream/keep-old", True, True)
