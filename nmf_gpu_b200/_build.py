"""In-tree build of libnmf_b200.so (hand-written sm_100a kernels + the C ABI of include/nmf.h).

nvcc only, no torch, no CMake: one object per source (compiled in parallel, rebuilt when the source
or any header is newer), then one link.  cudart is linked statically and neither libcuda nor NCCL is
linked at all (tensor-map encoding goes through cudaGetDriverEntryPoint, NCCL is dlopen'ed), so the
library loads on a machine without a GPU driver — its entry points then fail loudly.
"""
import os
import shutil
import subprocess
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libnmf_b200.so")
SOURCES = ["api.cpp", "dist.cpp", "session.cu", "kernels_small.cu", "kernels_simt.cu", "kernels_umma.cu", "kernels_fused.cu"]
HEADERS = ["common.hpp", "session.hpp", "umma_ptx.cuh", "tf32.cuh", "umma_host.inc", "umma_fused.cuh", "umma_common.cuh", os.path.join("..", "..", "include", "nmf.h")]
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]


def nvcc_path():
    for c in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", shutil.which("nvcc")):
        if c and os.path.exists(c):
            return c
    raise RuntimeError("nvcc not found")


def _newest_header():
    return max(os.path.getmtime(os.path.join(CSRC, f)) for f in HEADERS)


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(os.path.join(CSRC, f)) > t for f in SOURCES + HEADERS)


def build(force=False, verbose=False):
    if not force and not needs_build():
        return LIB
    os.makedirs(OBJ, exist_ok=True)
    nvcc, hdr = nvcc_path(), _newest_header()
    flags = ARCH + ["-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC", "-ccbin", "/usr/bin/g++"]
    if os.environ.get("NMF_BUILD_DIAG"):          # timing experiments that give wrong results (kernels_umma.cu): never in a release build
        flags += ["-DNMF_DIAG=1"]
    if verbose:
        flags += ["-Xptxas=-v"]

    def compile_one(src):
        s, o = os.path.join(CSRC, src), os.path.join(OBJ, src + ".o")
        if not force and os.path.exists(o) and os.path.getmtime(o) > max(os.path.getmtime(s), hdr):
            return o
        cmd = [nvcc] + flags + ["-x", "cu", "-c", s, "-o", o]
        if verbose:
            print(" ".join(cmd))
        subprocess.check_call(cmd)
        return o

    with ThreadPoolExecutor(max_workers=len(SOURCES)) as ex:
        objs = list(ex.map(compile_one, SOURCES))
    cmd = [nvcc] + ARCH + ["-shared", "-cudart", "static", "-ccbin", "/usr/bin/g++", "-o", LIB] + objs + ["-ldl", "-lpthread", "-lrt"]
    subprocess.check_call(cmd)
    return LIB


if __name__ == "__main__":
    import sys
    build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(LIB)
