"""In-tree build of libnmf_b200.so (hand-written sm_100a kernels + the C ABI of include/nmf.h).

One nvcc invocation, no torch, no CMake.  cudart is linked statically and neither libcuda nor NCCL
is linked at all (tensor-map encoding goes through cudaGetDriverEntryPoint, NCCL is dlopen'ed), so
the library loads on a machine without a GPU driver — its entry points then fail loudly.
"""
import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libnmf_b200.so")
SOURCES = ["api.cpp", "dist.cpp", "session.cu", "kernels_small.cu", "kernels_simt.cu", "kernels_umma.cu"]
HEADERS = ["common.hpp", "session.hpp", "umma_ptx.cuh", "tf32.cuh", "umma_host.inc", "umma_fused.inc", os.path.join("..", "..", "include", "nmf.h")]


def nvcc_path():
    for c in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", shutil.which("nvcc")):
        if c and os.path.exists(c):
            return c
    raise RuntimeError("nvcc not found")


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(os.path.join(CSRC, f)) > t for f in SOURCES + HEADERS)


def build(force=False, verbose=False):
    if not force and not needs_build():
        return LIB
    cmd = [nvcc_path(), "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
           "-Xcompiler", "-fPIC", "-shared", "-cudart", "static", "-ccbin", "/usr/bin/g++",
           "-x", "cu", "-o", LIB] + [os.path.join(CSRC, f) for f in SOURCES] + ["-ldl", "-lpthread", "-lrt"]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
        print(" ".join(cmd))
    subprocess.check_call(cmd)
    return LIB


if __name__ == "__main__":
    import sys
    build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(LIB)
