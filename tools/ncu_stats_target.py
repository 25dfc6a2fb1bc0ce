#!/usr/bin/env python
"""Short program for an ncu launch list of the display statistics on a 4096^2 image (vmed + vmad chain)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import aflak_b200 as afk

afk.set_device(0)
rng = np.random.default_rng(4)
img = afk.Image.upload(np.exp(rng.standard_normal((4096, 4096))).astype(np.float32))
for _ in range(3):
    print(img.vmed_vmad())
