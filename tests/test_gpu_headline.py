"""GPU parity tests in the regime of the headline benchmark (BASELINE configs C2-C5), all against the fp64 oracle
through the C ABI: long reductions (M = 16384: 64 row tiles, 512 TMEM promotions per work item; N = 65536: the
split-N fold), K = 512, the full ISMIR problem in exact semantics, sampled columns / rows of the FULL-size C4 problem
(indices past 2^32 elements), the plain `update_div` symbol and the `examples/nmf` binary.

Tolerances as tests/test_gpu_parity.py: half-steps 2e-5 (max |d| / max(|ref|, 1e-6 max|ref|)); 100 iterations 1e-4 on
W and H, 1e-5 relative on the KL divergence (BASELINE.json north_star).  The measured margins are printed (-s) and
written to gpurun_out/headline_margins.jsonl when that directory exists.
"""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from util import fro_rel, max_rel, rand_problem, survivors

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN_SURVIVORS = [1, 22, 30, 34, 43, 45, 57, 70, 88, 89, 93, 96, 110, 113, 118, 121, 126]


@pytest.fixture(scope="module")
def api():
    import nmf_gpu_b200.api as a
    a.lib()
    return a


@pytest.fixture(scope="module")
def oracle():
    from oracle import oracle as o
    o.lib()
    return o


def record(name, **kw):
    line = json.dumps(dict(test=name, **kw))
    print("MARGIN " + line)
    out = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(out):
        with open(os.path.join(out, "headline_margins.jsonl"), "a") as f:
            f.write(line + "\n")


# ---------------------------------------------------------------------------------------------
# (a) fused half-steps at C4-slice sizes
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("shape", [(16384, 2048, 256), (2048, 65536, 256)])
def test_fused_half_steps_c4_slices(api, oracle, shape):
    """16384 x 2048: the H step reduces over M = 16384 (64 row tiles; the rows of W split over work items, partial sums
    folded by k_update_h) and the W step runs 64 lane-row blocks.  2048 x 65536: the W step reduces over 65536 columns
    (split over work items, ordered fold in k_update_w); the H step's 256 column blocks run whole waves unsplit."""
    M, N, K = shape
    W, H, X = rand_problem(M, N, K, seed=M % 1000 + 3)
    with api.Session(M, N, K) as s:
        assert s.fused_steps() == 3
    H1 = api.dbg_h_step(W, H, X, api.options())
    eh = max_rel(H1, oracle.h_step(W, H, X, "exact", acc64=True))
    W1 = api.dbg_w_step(W, H, X, api.options())
    ew = max_rel(W1, oracle.w_step(W, H, X, "exact", acc64=True))
    record("fused_half_steps_c4_slices", shape=shape, H=eh, W=ew, tol=2e-5)
    assert eh < 2e-5 and ew < 2e-5


# ---------------------------------------------------------------------------------------------
# (b) 100 iterations at the C4 slice: the north-star tolerance where the reductions are longest
# ---------------------------------------------------------------------------------------------
def test_c4_slice_100_iterations(api, oracle):
    M, N, K = 16384, 2048, 256
    W, H, X = rand_problem(M, N, K, seed=4)
    Wo, Ho, _, cost = oracle.update_div(W, H, X, 0.0, 100, "exact", acc64=True)
    Wd, Hd = W.copy(order="F"), H.copy(order="F")
    res = api.update_div(Wd, Hd, X, 0.0, 100)
    assert res.iterations == 100
    ew, eh = max_rel(Wd, Wo), max_rel(Hd, Ho)
    ek = abs(oracle.kl(Wd, Hd, X) - cost) / cost
    record("c4_slice_100_iterations", shape=(M, N, K), W=ew, H=eh, KL=ek, tol=(1e-4, 1e-5))
    assert ew <= 1e-4 and eh <= 1e-4 and ek <= 1e-5


# ---------------------------------------------------------------------------------------------
# (c) K = 512 (BASELINE C5): the two-kernel path (P wider than the free TMEM columns)
# ---------------------------------------------------------------------------------------------
def test_k512_half_steps_and_iterations(api, oracle):
    M, N, K = 8192, 1024, 512
    W, H, X = rand_problem(M, N, K, seed=512)
    H1 = api.dbg_h_step(W, H, X, api.options())
    eh = max_rel(H1, oracle.h_step(W, H, X, "exact", acc64=True))
    W1 = api.dbg_w_step(W, H, X, api.options())
    ew = max_rel(W1, oracle.w_step(W, H, X, "exact", acc64=True))
    assert eh < 2e-5 and ew < 2e-5
    Wo, Ho, _, cost = oracle.update_div(W, H, X, 0.0, 30, "exact", acc64=True)
    Wd, Hd = W.copy(order="F"), H.copy(order="F")
    res = api.update_div(Wd, Hd, X, 1e-12, 30)                        # thresh > 0: one cost check at iteration 26
    assert res.iterations == 30
    iw, ih = max_rel(Wd, Wo), max_rel(Hd, Ho)
    ek = abs(oracle.kl(Wd, Hd, X) - cost) / cost
    record("k512", shape=(M, N, K), H_step=eh, W_step=ew, W30=iw, H30=ih, KL=ek)
    assert iw <= 1e-4 and ih <= 1e-4 and ek <= 1e-5


@pytest.mark.parametrize("K", [80, 96, 200])
def test_k_not_multiple_of_64_runs_fused(api, oracle, K):
    """exact semantics pad K to a multiple of 64 (zero fill: zero rows / columns are fixed points of the update) when that
    is what lets both half-iterations fuse; NMF_PAD_K64=0 keeps the 32-padding and the two-kernel path — same results"""
    M, N = 700, 1500
    W, H, X = rand_problem(M, N, K, seed=K)
    with api.Session(M, N, K) as s:
        assert s.fused_steps() == 3
    Wo, Ho, _, cost = oracle.update_div(W, H, X, 0.0, 30, "exact", acc64=True)
    Wd, Hd = W.copy(order="F"), H.copy(order="F")
    res = api.update_div(Wd, Hd, X, 1e-12, 30)
    assert res.iterations == 30
    assert max_rel(Wd, Wo) <= 1e-4 and max_rel(Hd, Ho) <= 1e-4
    assert abs(res.cost - oracle.update_div(W, H, X, 0.0, 25, "exact", acc64=True)[3]) / cost < 1e-4
    os.environ["NMF_PAD_K64"] = "0"
    try:
        with api.Session(M, N, K) as s:
            assert s.fused_steps() == 0
        W2, H2 = W.copy(order="F"), H.copy(order="F")
        api.update_div(W2, H2, X, 1e-12, 30)
    finally:
        del os.environ["NMF_PAD_K64"]
    assert max_rel(W2, Wo) <= 1e-4 and max_rel(H2, Ho) <= 1e-4


@pytest.mark.parametrize("shape", [(513, 3000, 40), (257, 700, 64), (514, 1500, 128), (1027, 900, 192), (516, 2100, 256), (2049, 1300, 128)])
def test_w_step_tail_rows(api, oracle, shape, monkeypatch):
    """spectrogram heights 2^k + 1 (C2: 513, C3: 2049): the 1..4 rows beyond the last whole 256-row block are taken out of
    the fused W kernel (they would cost a whole 256-row work item per column split) and run on the CUDA cores
    (launch_w_tail; on from 65536 columns, where it pays — NMF_W_TAIL=2 forces it here, =0 keeps the rows in the fused
    kernel).  Both ways against the oracle, half-step and 30 iterations."""
    M, N, K = shape
    W, H, X = rand_problem(M, N, K, seed=M + K)
    ref = oracle.w_step(W, H, X, "exact", acc64=True)
    Wo, Ho, _, cost = oracle.update_div(W, H, X, 0.0, 30, "exact", acc64=True)
    for tail in ("2", "0"):
        monkeypatch.setenv("NMF_W_TAIL", tail)
        W1 = api.dbg_w_step(W, H, X, api.options())
        assert max_rel(W1, ref) < 2e-5, f"NMF_W_TAIL={tail}"
        assert max_rel(W1[256 * (M // 256):], ref[256 * (M // 256):]) < 2e-5        # the tail rows themselves
        Wd, Hd = W.copy(order="F"), H.copy(order="F")
        res = api.update_div(Wd, Hd, X, 1e-12, 30)
        assert res.iterations == 30
        assert max_rel(Wd, Wo) <= 1e-4 and max_rel(Hd, Ho) <= 1e-4
        assert abs(oracle.kl(Wd, Hd, X) - cost) / cost <= 1e-5


# ---------------------------------------------------------------------------------------------
# (d) BASELINE C2 as specified, exact semantics: 513 x 20000, K = 40, seed 1, 200 iterations
# ---------------------------------------------------------------------------------------------
def test_c2_full_size_exact_200_iterations(api, oracle):
    M, N, K = 513, 20000, 40
    W, H, X = rand_problem(M, N, K, seed=1)
    Wo, Ho, _, cost = oracle.update_div(W, H, X, 0.0, 200, "exact", acc64=True)
    Wd, Hd = W.copy(order="F"), H.copy(order="F")
    res = api.update_div(Wd, Hd, X, 0.0, 200)
    assert res.iterations == 200
    ew, eh = max_rel(Wd, Wo), max_rel(Hd, Ho)
    ek = abs(oracle.kl(Wd, Hd, X) - cost) / cost
    fw, fh = fro_rel(Wd, Wo), fro_rel(Hd, Ho)
    record("c2_full_exact_200", W=ew, H=eh, KL=ek, fro_W=fw, fro_H=fh)
    # 200 iterations is twice the north-star horizon: the elementwise budget is the 100-iteration one times two,
    # the divergence and the norm-wise errors keep their bounds
    assert ek <= 1e-5 and fw <= 2e-5 and fh <= 2e-5
    assert ew <= 2e-4 and eh <= 2e-4


# ---------------------------------------------------------------------------------------------
# (e) the FULL-size C4 problem, spot-checked: one H step on sampled columns, one W step on sampled rows
# ---------------------------------------------------------------------------------------------
def test_c4_full_size_spot_check(api, oracle):
    """16384 x 262144, K = 256 generated on the device (X alone is 17.2 GB: element indices pass 2^32).  The H step is
    column-local, so 64 sampled columns (the first, the last, both sides of the 2^32-element boundary at column 131072 and
    of 256-column block edges, and random ones) are checked against oracle.h_step on those columns; the W step is
    row-local, so 24 sampled rows are checked against oracle.w_step on those rows and ALL columns."""
    M, N, K = 16384, 262144, 256
    free = 0
    try:
        import torch
        free = torch.cuda.mem_get_info()[0]
    except Exception:
        pass
    if free and free < 30e9:
        pytest.skip("needs ~19 GB of free HBM")
    rng = np.random.RandomState(7)
    cols = sorted(set([0, 1, 255, 256, 131071, 131072, 131073, N - 257, N - 256, N - 1] +
                      rng.randint(0, N, size=54).tolist()))
    rows = sorted(set([0, 255, 256, 8191, 8192, M - 1] + rng.randint(0, M, size=18).tolist()))
    with api.Session(M, N, K) as s:
        assert s.fused_steps() == 3
        s.fill_uniform(4)
        Wfull, Hfull0 = s.download()
        Xc, Hc0 = s.download_cols(cols)
        assert np.array_equal(Hc0, Hfull0[:, cols])
        del Hfull0
        s.half_step("h")
        _, Hc1 = s.download_cols(cols)
        Href = oracle.h_step(Wfull, Hc0, Xc, "exact", acc64=True)
        eh = max_rel(Hc1, Href)
        # W step from the H the device now holds
        _, Hfull1 = s.download()
        Xr, Wr0 = s.download_rows(rows)
        assert np.array_equal(Wr0, Wfull[rows])
        s.half_step("w")
        _, Wr1 = s.download_rows(rows)
        Wref = oracle.w_step(Wr0, Hfull1, Xr, "exact", acc64=True)
        ew = max_rel(Wr1, Wref)
    record("c4_full_size_spot_check", cols=len(cols), rows=len(rows), H=eh, W=ew, tol=2e-5)
    assert eh < 2e-5 and ew < 2e-5


def test_c3_shape_slice_iterations(api, oracle):
    """BASELINE C3's shape family (2049 rows: eight 256-row blocks + ONE row; K = 128: the 4-worker-warp fused kernels
    with double-buffered S tiles) on a 4096-column slice of the planted target: 50 iterations with a cost check"""
    M, N, K = 2049, 4096, 128
    rng = np.random.RandomState(3)
    X = np.asfortranarray(((rng.rand(M, K) @ rng.rand(K, N)) * (1 + 0.05 * (2 * rng.rand(M, N) - 1))).astype(np.float32))
    W, H, _ = rand_problem(M, N, K, seed=33)
    Wo, Ho, _, cost = oracle.update_div(W, H, X, 0.0, 50, "exact", acc64=True)
    Wd, Hd = W.copy(order="F"), H.copy(order="F")
    res = api.update_div(Wd, Hd, X, 1e-12, 50)
    assert res.iterations == 50
    ew, eh = max_rel(Wd, Wo), max_rel(Hd, Ho)
    ek = abs(oracle.kl(Wd, Hd, X) - cost) / cost
    record("c3_shape_slice_50", W=ew, H=eh, KL=ek)
    assert ew <= 1e-4 and eh <= 1e-4 and ek <= 1e-5


# ---------------------------------------------------------------------------------------------
# (f) the boundary itself: the plain update_div symbol, and the examples/nmf binary on the reference's file flow
# ---------------------------------------------------------------------------------------------
def test_plain_update_div_symbol(api, oracle, c1, monkeypatch):
    """`void update_div(matrix W, matrix H, matrix X, float, int, double t[10], int)` — structs by value, semantics from
    the environment as for a C caller — against the golden (refcompat, 200 iterations, cuda/nmf.cu:10-11)"""
    monkeypatch.setenv("NMF_SEMANTICS", "refcompat")
    W = c1["W"].copy(order="F"); H = c1["H"].copy(order="F")
    t = np.zeros(10)
    api.update_div_plain(W, H, c1["X"], 0.0, 200, t=t, verbose=0)
    gold = oracle.kl(c1["Wtest"], c1["Htest"], c1["X"])
    assert abs(oracle.kl(W, H, c1["X"]) - gold) / gold <= 1e-5
    assert survivors(W) == GOLDEN_SURVIVORS
    assert fro_rel(W, c1["Wtest"]) <= 2e-3 and fro_rel(H, c1["Htest"]) <= 2e-3
    assert t[9] > 0 and t[3] > 0
    monkeypatch.setenv("NMF_SEMANTICS", "exact")
    W = c1["W"].copy(order="F"); H = c1["H"].copy(order="F")
    api.update_div_plain(W, H, c1["X"], 0.0, 100)
    assert abs(oracle.kl(W, H, c1["X"]) - 105982.05) / 105982.05 < 1e-5     # BASELINE.md known answer


def test_examples_nmf_binary_reference_flow(api, c1, tmp_path):
    """the reference's CLI flow (cuda/nmf.cu:37-45: ../X.bin ../H.bin ../W.bin -> ../Wout.bin ../Hout.bin, 200 iterations)
    with the files examples/matrix_export.py writes, through examples/nmf, checked the way test_output.sh is meant
    (examples/compare_output.py: tolerance instead of md5) against the reference's Wtest.bin / Htest.bin"""
    exe = os.path.join(ROOT, "examples", "nmf")
    if not os.path.exists(exe):
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "examples")])
    subprocess.check_call([sys.executable, os.path.join(ROOT, "examples", "matrix_export.py"), "--config", "c1", "--out", str(tmp_path)])
    cwd = tmp_path / "cuda"; cwd.mkdir()
    env = dict(os.environ, NMF_SEMANTICS="refcompat")
    r = subprocess.run([exe], cwd=cwd, env=env, capture_output=True, text=True, timeout=600)      # no arguments: the reference's defaults
    assert r.returncode == 0, r.stderr + r.stdout
    assert "read ../X.bin [4096x352]" in r.stdout and "write ../Wout.bin [4096x128]" in r.stdout   # nmf.cu:215,258
    api.write_matrix(c1["Wtest"], str(tmp_path / "Wtest.bin")); api.write_matrix(c1["Htest"], str(tmp_path / "Htest.bin"))
    r2 = subprocess.run([sys.executable, os.path.join(ROOT, "examples", "compare_output.py"), str(tmp_path)],
                        capture_output=True, text=True)
    assert r2.returncode == 0, r2.stdout + r2.stderr
    # argv form: directory, iterations, threshold, verbose, GPUs
    r3 = subprocess.run([exe, str(tmp_path), "50", "0.001", "1", "1"], cwd=cwd, env=env, capture_output=True, text=True, timeout=600)
    assert r3.returncode == 0 and "iter 25  cost" in r3.stdout and "seconds:" in r3.stdout, r3.stdout + r3.stderr
