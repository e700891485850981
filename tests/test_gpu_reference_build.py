"""Parity against the reference's OWN CUDA program (oracle/_ref/nmf_ref, built by oracle/Makefile
from /root/reference/cuda/*.cu, unmodified) run on this GPU: BASELINE config C1, the only
configuration the unmodified reference can run.  Its flow (cuda/nmf.cu:37-45): read
../X.bin ../H.bin ../W.bin, 200 iterations, write ../Wout.bin ../Hout.bin."""
import os
import struct
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from util import fro_rel, survivors

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "nmf_ref")


def run_reference(tmp_path):
    """matrix_export.py's files, then the reference binary from a sub-directory (it uses ../)."""
    np.random.seed(0)
    X = np.random.rand(4096, 350).astype(np.float32)
    W = np.random.rand(4096, 128).astype(np.float32)
    H = np.random.rand(128, 350).astype(np.float32)
    for name, a in (("X", X), ("W", W), ("H", H)):
        with open(tmp_path / f"{name}.bin", "wb") as f:
            f.write(struct.pack("ii", *a.shape)); f.write(a.tobytes())
    cwd = tmp_path / "cuda"; cwd.mkdir()
    env = dict(os.environ, NVIDIA_TF32_OVERRIDE="0")
    env["LD_LIBRARY_PATH"] = "/usr/local/cuda/lib64:" + env.get("LD_LIBRARY_PATH", "")
    r = subprocess.run([REF_BIN], cwd=cwd, env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr + r.stdout
    return r.stdout


@pytest.mark.skipif(not os.path.exists(REF_BIN), reason="oracle/_ref/nmf_ref not built (make -C oracle ref)")
def test_refcompat_matches_reference_cuda_build(tmp_path, c1):
    import nmf_gpu_b200.api as api
    from oracle import oracle
    run_reference(tmp_path)
    Wr = api.read_matrix(str(tmp_path / "Wout.bin")); Hr = api.read_matrix(str(tmp_path / "Hout.bin"))
    out = os.path.join(ROOT, "gpurun_out")      # keep the reference's output on this hardware as an artefact
    if os.path.isdir(out):
        np.savez_compressed(os.path.join(out, "ref_b200_c1.npz"), Wout=Wr, Hout=Hr)
    # (a) the reference on this GPU still reproduces its own golden (tolerance-level; md5 is GPU-specific)
    gold = oracle.kl(c1["Wtest"], c1["Htest"], c1["X"])
    kl_ref = oracle.kl(Wr, Hr, c1["X"])
    assert abs(kl_ref - gold) / gold <= 1e-5 and survivors(Wr) == survivors(c1["Wtest"])
    # (b) our refcompat path against the reference build run here
    W = c1["W"].copy(order="F"); H = c1["H"].copy(order="F")
    api.update_div(W, H, c1["X"], 0.0, 200, opt=api.options(semantics="refcompat"))
    assert abs(oracle.kl(W, H, c1["X"]) - kl_ref) / kl_ref <= 1e-5
    assert survivors(W) == survivors(Wr)
    assert fro_rel(W, Wr) <= 2e-3 and fro_rel(H, Hr) <= 2e-3
