"""CPU: the bracket test of the one-pass select survives compilation (aflak_b200/csrc/stats.cu, stat_bracket_pass_kernel).

Round 2 met a build in which nvcc 12.9 emitted one inlined copy of the pass's classification WITHOUT its `key > hi` term (two
instantiations of the stream inlined side by side in one kernel): the histogram reduction then addressed shared memory with an
unbounded bin and the kernel faulted on the device.  The code was restructured; this test compiles the translation unit to PTX
(nvcc cross-compiles without a GPU) and checks, for every inlined copy in both instantiations, that the shared-memory reduction
is reached only through the branch on `key - lo > hi - lo`, against one and the same span `hi - lo`.
"""
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _nvcc():
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", shutil.which("nvcc")):
        if cand and os.path.exists(cand):
            return cand
    return None


def test_every_copy_of_the_bracket_test_guards_its_histogram_reduction(tmp_path):
    nvcc = _nvcc()
    if not nvcc:
        pytest.skip("no nvcc on this machine")
    ptx = tmp_path / "stats.ptx"
    env = dict(os.environ)
    env.pop("CC", None)
    env.pop("CXX", None)
    subprocess.check_call([nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-fmad=false", "-I", os.path.join(ROOT, "include"),
                           "-ptx", os.path.join(ROOT, "aflak_b200", "csrc", "stats.cu"), "-o", str(ptx)], env=env)
    txt = ptx.read_text()
    for tag in ("ILb0", "ILb1"):  # normalised values, raw keys
        m = re.search(r"\.entry _ZN3afk24stat_bracket_pass_kernel" + tag + r".*?\n\}\n", txt, re.S)
        assert m, f"stat_bracket_pass_kernel<{tag}> not found in the PTX"
        body = m.group(0).splitlines()
        reds = [i for i, line in enumerate(body) if "red.shared.add.u32" in line]
        assert len(reds) >= 21, f"{tag}: expected at least the 16 + 4 + 1 inlined copies, found {len(reds)}"
        spans = set()
        for i in reds:
            branch = next((j for j in range(i - 1, max(i - 14, 0), -1) if re.match(r"\s*@!?%p\d+\s+bra", body[j])), None)
            assert branch is not None, f"{tag}: reduction at PTX line {i} is not behind a branch"
            pred = re.match(r"\s*@!?(%p\d+)\s+bra", body[branch]).group(1)
            setp = next((body[k] for k in range(branch - 1, max(branch - 12, 0), -1)
                         if re.search(r"setp\.\w+\.u32\s+" + re.escape(pred) + ",", body[k])), None)
            assert setp, f"{tag}: no unsigned compare defines the guard of the reduction at PTX line {i}"
            mm = re.search(r"setp\.(gt|le)\.u32\s+%p\d+, (%r\d+), (%r\d+);", setp)
            assert mm, f"{tag}: the guard of the reduction at PTX line {i} is not the span compare: {setp.strip()}"
            spans.add(mm.group(3))
        # one span: either one register, or re-materialised copies of the same `hi - lo` (the compiler recomputes it under
        # register pressure)
        defs = set()
        for reg in spans:
            d = [re.search(r"sub\.s32\s+" + re.escape(reg) + r", (%r\d+), (%r\d+);", line) for line in body]
            d = [x.groups() for x in d if x]
            assert d, f"{tag}: span register {reg} is not defined as a difference"
            defs.update(d)
        assert len(defs) == 1, f"{tag}: the copies compare against different spans {defs}"
