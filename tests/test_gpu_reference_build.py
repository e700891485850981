"""Parity against the reference's OWN CUDA program (oracle/_ref/nmf_ref, built by oracle/Makefile
from /root/reference/cuda/*.cu, unmodified) run on this GPU: BASELINE config C1, the only
configuration the unmodified reference can run, and C2 (513x20000, K=40) with the same sources
compiled -DNDEBUG (oracle/_ref/nmf_ref_ndebug: the inverted assert at cuda/matrix.cu:287,406 aborts
when K_p < 128).  Its flow (cuda/nmf.cu:37-45): read ../X.bin ../H.bin ../W.bin,
200 iterations, write ../Wout.bin ../Hout.bin."""
import json
import os
import struct
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from util import fro_rel, survivors

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "nmf_ref")


REF_BIN_NDEBUG = os.path.join(ROOT, "oracle", "_ref", "nmf_ref_ndebug")


def run_reference(tmp_path, binary=REF_BIN, M=4096, N=350, K=128, seed=0):
    """matrix_export.py's files (its sizes and seed by default), then the reference binary from a
    sub-directory (it uses ../).  Returns the matrices as read_matrix loads them (payload column-major)."""
    np.random.seed(seed)
    X = np.random.rand(M, N).astype(np.float32)
    W = np.random.rand(M, K).astype(np.float32)
    H = np.random.rand(K, N).astype(np.float32)
    for name, a in (("X", X), ("W", W), ("H", H)):
        with open(tmp_path / f"{name}.bin", "wb") as f:
            f.write(struct.pack("ii", *a.shape)); f.write(a.tobytes())
    cwd = tmp_path / "cuda"; cwd.mkdir()
    env = dict(os.environ, NVIDIA_TF32_OVERRIDE="0")
    env["LD_LIBRARY_PATH"] = "/usr/local/cuda/lib64:" + env.get("LD_LIBRARY_PATH", "")
    r = subprocess.run([binary], cwd=cwd, env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr + r.stdout
    as_loaded = lambda a: a.reshape(-1).reshape(a.shape, order="F").copy(order="F")
    return as_loaded(W), as_loaded(H), as_loaded(X)


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


@pytest.mark.skipif(not os.path.exists(REF_BIN_NDEBUG), reason="oracle/_ref/nmf_ref_ndebug not built (make -C oracle ref)")
def test_refcompat_matches_reference_cuda_build_c2(tmp_path):
    """BASELINE config C2 (ISMIR shape 513x20000, K=40 -> padded 544x20000, K_p=64; seed 1, 200 iterations):
    here rows_padded <= 1024, so the reference's row_div DOES run (with its subsampled row sums of H,
    cuda/matrix.cu:212-217,724-731) — the branch of refcompat that C1 never takes."""
    import nmf_gpu_b200.api as api
    from oracle import oracle
    W0, H0, X = run_reference(tmp_path, REF_BIN_NDEBUG, 513, 20000, 40, seed=1)
    Wr = api.read_matrix(str(tmp_path / "Wout.bin")); Hr = api.read_matrix(str(tmp_path / "Hout.bin"))
    assert Wr.shape == (513, 40) and Hr.shape == (40, 20000)
    assert np.isfinite(Wr).all() and np.isfinite(Hr).all()
    kl_ref = oracle.kl(Wr, Hr, X)
    W = W0.copy(order="F"); H = H0.copy(order="F")
    api.update_div(W, H, X, 0.0, 200, opt=api.options(semantics="refcompat"))
    kl = oracle.kl(W, H, X)
    print(f"C2 reference CUDA build KL {kl_ref:.6f}, ours (refcompat) {kl:.6f}, "
          f"Frobenius W {fro_rel(W, Wr):.2e} H {fro_rel(H, Hr):.2e}, survivors {len(survivors(Wr))}")
    # 36 of the 40 components collapse on this input and the trajectory is sensitive to rounding: the fp32 and fp64
    # oracles differ by 8.8e-4 in KL (Frobenius W 1.9e-3) after 200 iterations.  Measured on B200: reference build vs
    # ours 7.4e-5 in KL, Frobenius W 1.6e-4 / H 2.8e-5 (the reference build sits 6.6e-6 from the fp64 oracle).
    assert abs(kl - kl_ref) / kl_ref <= 5e-4
    assert survivors(W) == survivors(Wr)
    assert fro_rel(W, Wr) <= 2e-3 and fro_rel(H, Hr) <= 2e-3


REF_PATCHED = os.path.join(ROOT, "baseline", "_ref", "nmf_ref_patched")


def run_patched(tmp_path, M, N, K, seed, iters):
    """the patched reference build (baseline/make_ref_patched.py: 64-bit sizes, large grids, TRUE sums, row_div always
    runs — i.e. the reference's CUDA code computing what README / paper specify) on matrix_export-style files"""
    np.random.seed(seed)
    X = np.random.rand(M, N).astype(np.float32)
    W = np.random.rand(M, K).astype(np.float32)
    H = np.random.rand(K, N).astype(np.float32)
    for name, a in (("X", X), ("W", W), ("H", H)):
        with open(tmp_path / f"{name}.bin", "wb") as f:
            f.write(struct.pack("ii", *a.shape)); f.write(a.tobytes())
    env = dict(os.environ, NVIDIA_TF32_OVERRIDE="0")
    env["LD_LIBRARY_PATH"] = "/usr/local/cuda/lib64:" + env.get("LD_LIBRARY_PATH", "")
    r = subprocess.run([REF_PATCHED, str(M), str(N), str(K), str(iters), str(tmp_path)], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr + r.stdout
    as_loaded = lambda a: a.reshape(-1).reshape(a.shape, order="F").copy(order="F")
    return as_loaded(W), as_loaded(H), as_loaded(X)


@pytest.mark.skipif(not os.path.exists(REF_PATCHED), reason="baseline/_ref/nmf_ref_patched not built (python baseline/make_ref_patched.py)")
@pytest.mark.parametrize("cfg", [("c1", 4096, 350, 128, 0), ("c2", 513, 20000, 40, 1)])
def test_exact_matches_patched_reference_cuda_build(tmp_path, cfg):
    """exact semantics against the reference's OWN CUDA code (cublasSgemm + its kernels, with the sums repaired) on this
    GPU, BASELINE C1 and C2 at full size, 100 iterations: the north-star tolerance between the two CUDA programs, and each
    of them against the fp64 oracle — which pins the oracle's `exact` mode to the reference's arithmetic, not only to the
    paper."""
    import nmf_gpu_b200.api as api
    from oracle import oracle
    from util import max_rel
    name, M, N, K, seed = cfg
    W0, H0, X = run_patched(tmp_path, M, N, K, seed, 100)
    Wr = api.read_matrix(str(tmp_path / "Wout.bin")); Hr = api.read_matrix(str(tmp_path / "Hout.bin"))
    W = W0.copy(order="F"); H = H0.copy(order="F")
    api.update_div(W, H, X, 0.0, 100)
    Wo, Ho, _, cost = oracle.update_div(W0, H0, X, 0.0, 100, "exact", acc64=True)
    kl_ref, kl = oracle.kl(Wr, Hr, X), oracle.kl(W, H, X)
    m = dict(ours_vs_ref=(max_rel(W, Wr), max_rel(H, Hr)), ref_vs_oracle=(max_rel(Wr, Wo), max_rel(Hr, Ho)),
             ours_vs_oracle=(max_rel(W, Wo), max_rel(H, Ho)), kl=(abs(kl - kl_ref) / kl_ref, abs(kl_ref - cost) / cost))
    print(f"{name} exact, 100 iterations: {m}")
    out = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(out):
        with open(os.path.join(out, "patched_reference_parity.jsonl"), "a") as f:
            f.write(json.dumps(dict(config=name, **m)) + "\n")
    # the reference accumulates in plain fp32 (cublasSgemm): it sits as far from the fp64 oracle as our 3xTF32 path does,
    # so the two CUDA programs may differ by up to the sum of both distances
    assert m["kl"][0] <= 1e-5 and m["kl"][1] <= 1e-5
    assert max(m["ref_vs_oracle"]) <= 1e-4 and max(m["ours_vs_oracle"]) <= 1e-4
    assert max(m["ours_vs_ref"]) <= 2e-4
