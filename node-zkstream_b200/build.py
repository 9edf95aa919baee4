"""Builds libzkbatch.so (the CUDA codec + C ABI) in-tree with nvcc for sm_100a.

    python node-zkstream_b200/build.py [--force]

The shared library lands next to this file so it travels with the repository
snapshot to the GPU box; there is no JIT cache and no pip install involved.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libzkbatch.so")
SOURCES = ["zk_ctx.cu", "zk_decode.cu", "zk_encode.cu", "zk_stats.cu"]
HEADERS = [os.path.join(CSRC, "zk_internal.h"), os.path.join(CSRC, "zk_decode_direct.cuh"),
           os.path.join(HERE, "..", "include", "zkbatch.h")]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC,-fvisibility=hidden",
    "--cudart", "shared",
]


def _stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, s) for s in SOURCES] + HEADERS + [os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps)


GEN_LIB = os.path.join(HERE, "libzkgen.so")


def build_gen(force=False):
    """libzkgen.so: the host-side synthetic traffic generator (plain C, no codec in it)."""
    src = os.path.join(CSRC, "zkgen.c")
    if force or not os.path.exists(GEN_LIB) or os.path.getmtime(GEN_LIB) < os.path.getmtime(src):
        subprocess.run(["gcc", "-O2", "-Wall", "-fPIC", "-shared", "-o", GEN_LIB, src, "-lpthread"], check=True)
    return GEN_LIB


def build(force=False, verbose=False):
    build_gen(force)
    if not force and not _stale():
        return LIB
    objs = []
    for s in SOURCES:
        o = os.path.join(CSRC, s.replace(".cu", ".o"))
        cmd = [NVCC] + FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", os.path.join(CSRC, s), "-o", o]
        subprocess.run(cmd, check=True)
        objs.append(o)
    subprocess.run([NVCC, "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "--cudart", "shared", "-o", LIB] + objs,
                   check=True)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
