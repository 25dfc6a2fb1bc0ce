"""Build libaflak_b200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

    python aflak_b200/build.py [--force] [-v]

Staleness is decided by a content hash of the sources (kept next to the library in libaflak_b200.so.stamp), not by
mtimes: the built library travels to the GPU box inside a snapshot whose mtimes are arbitrary.
"""
import hashlib
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libaflak_b200.so")
STAMP = LIB + ".stamp"
SOURCES = ["kernels.cu", "noalloc.cu", "axis0.cu", "ingest.cu", "stats.cu", "spatial.cu", "fits.cpp", "stats.cpp", "ops_spatial.cpp", "runtime.cpp", "ops.cpp", "display.cpp", "graph.cpp", "shard.cpp"]
HEADERS = ["kernels.h", "host.h", os.path.join("..", "..", "include", "aflak_b200.h")]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-fmad=false",  # the reference never contracts a*b+c: elementwise results must be bit-identical
    "-Xcompiler", "-fPIC,-Wall,-Wextra",
    "-x", "cu",
    "-shared",
    "-cudart", "static",
    "-Xlinker", "--no-undefined",  # a missing definition fails the build, not the first dlopen on the GPU box
    "--threads", "0",  # compile the translation units in parallel
]


def _nvcc():
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    return "nvcc"


def source_hash():
    h = hashlib.sha256()
    h.update(" ".join(NVCC_FLAGS).encode())
    for rel in SOURCES + HEADERS:
        with open(os.path.join(CSRC, rel), "rb") as f:
            h.update(rel.encode())
            h.update(f.read())
    return h.hexdigest()


def needs_build():
    if not os.path.exists(LIB) or not os.path.exists(STAMP):
        return True
    with open(STAMP) as f:
        return f.read().strip() != source_hash()


def build(force=False, verbose=False):
    if not force and not needs_build():
        return LIB
    env = dict(os.environ)
    env.pop("CC", None)   # the image exports CC=/opt/gcc/bin/gcc; nvcc should use the gcc on PATH
    env.pop("CXX", None)
    cmd = [_nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB] + [os.path.join(CSRC, s) for s in SOURCES]
    subprocess.check_call(cmd, env=env)
    with open(STAMP, "w") as f:
        f.write(source_hash() + "\n")
    return LIB


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(LIB)
