"""Worker of test_median_one_pass_short_windows_subprocess: the one-pass median kernel on windows of 801..2100 frames
(the process is started with AFK_MEDIAN_SINGLE_PASS=2; MEDIAN_WORKER_NZ picks the window lengths, AFK_MEDIAN_Z the bracket's
half-width in sigmas: 0.5 makes most brackets miss or end on the wanted rank -- wing / open-bin re-collects and the even-count
successor above the bracket).  Exit status 0 = every pattern bit-exact against the oracle."""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)

import numpy as np

import aflak_b200 as gpu
import oracle as orc
from test_gpu_axis0 import _median_one_pass_patterns

orc.build()
gpu.set_device(0)
rng = np.random.default_rng(0xAF1A0044)
ny, nx = 97, 99
bad = 0
for nz in [int(x) for x in os.environ.get("MEDIAN_WORKER_NZ", "801,1300,1700,2100").split(",")]:
    for tag, host in _median_one_pass_patterns(rng, nz, ny, nx):
        cube = gpu.Image.upload(host)
        for start, end in [(0, nz), (0, nz - 1)] if nz > 801 else [(0, nz)]:
            got = gpu.call("median", cube, start, end).download()
            want = orc.median(host, start, end)
            if not np.array_equal(got.view(np.uint32), want.view(np.uint32)):
                bad += 1
                print(f"MISMATCH {tag} nz={nz} [{start},{end}): {int((got.view(np.uint32) != want.view(np.uint32)).sum())} spaxels")
print("mismatches:", bad)
sys.exit(1 if bad else 0)
