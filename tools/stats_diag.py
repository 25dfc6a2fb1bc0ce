#!/usr/bin/env python
"""vmed + vmad on one 4096^2 image, both select paths, with the stage trace (AFK_STATS_DEBUG=1) -- localises a failing launch."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import aflak_b200 as afk
import oracle

afk.set_device(0)
rng = np.random.default_rng(4)
a = np.exp(rng.standard_normal((2048, 2048))).astype(np.float32)
a[::1000, ::3] = np.nan
img = afk.Image.upload(a)
print("vmin/vmax", img.vmin_vmax(), flush=True)
for i in range(3):
    print("vmed/vmad", img.vmed_vmad(), flush=True)
print("oracle   ", oracle.vmed_vmad(a) if hasattr(oracle, "vmed_vmad") else "n/a", flush=True)
