"""Multi-GPU parity of the sharded path: the fused peer-store gather of the reduce kernel's epilogue, the
IPC + epoch-flag protocol of aflak_b200/csrc/shard.cpp, the min/max exchange and the RGBA band gather of the sharded
display path.  Needs two GPUs: skipped, with the reason shown, on a single-GPU box (bench.py --gpus N carries the same
check in its `gather_parity` block at every N the driver runs).
"""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
def test_two_rank_gather_parity(gpu):
    n = gpu.device_count()
    if n < 2:
        pytest.skip(f"needs 2 GPUs for the peer-store gather, this box has {n}; single-rank path: test_fused_gather_single_rank, "
                    "multi-rank: bench.py --gpus N gather_parity")
    world = 2
    env = dict(os.environ)
    env.setdefault("OMP_NUM_THREADS", "4")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", "29517", os.path.join(ROOT, "tests", "multirank_worker.py")]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=900, env=env, cwd=ROOT)
    out_dir = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(out_dir):  # evidence for the record: which memory paths ran (IPC peer stores, NVLink multicast)
        with open(os.path.join(out_dir, "multirank_worker_stdout.txt"), "w") as f:
            f.write(p.stdout[-8000:])
    assert p.returncode == 0 and f"MULTIRANK OK world={world}" in p.stdout, p.stdout[-4000:] + p.stderr[-4000:]
    assert "MULTICAST used" in p.stdout or "unavailable" in p.stdout or "no multicast mapping" in p.stdout, p.stdout[-2000:]
