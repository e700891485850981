"""Pins the CPU oracle (oracle/nmf_oracle.cpp) against the reference's own golden vectors and the
known-answer scalars of BASELINE.md.  CPU only."""
import numpy as np
import pytest

from oracle import oracle
from util import fro_rel, max_rel, survivors, rand_problem

GOLDEN_SURVIVORS = [1, 22, 30, 34, 43, 45, 57, 70, 88, 89, 93, 96, 110, 113, 118, 121, 126]


def test_eps_bits():
    # (float)2.2204E-16, cuda/matrix.cu:10 -> 0x257ffea4
    assert np.float32(oracle.lib().nmf_oracle_eps()).view(np.uint32) == 0x257FFEA4


def test_golden_is_what_survey_says(c1):
    assert c1["Wtest"].shape == (4096, 128) and c1["Htest"].shape == (128, 350)
    assert survivors(c1["Wtest"]) == GOLDEN_SURVIVORS
    kl = oracle.kl(c1["Wtest"], c1["Htest"], c1["X"])
    assert abs(kl - 4384096.64) / 4384096.64 < 1e-7
    assert abs(float(c1["X"].astype(np.float64).sum()) - 717076.78) < 0.5


def test_refcompat_reproduces_reference_golden(c1):
    """200 iterations, thresh 0 (cuda/nmf.cu:10-11), refcompat semantics == Wtest.bin / Htest.bin."""
    W, H, it, cost = oracle.update_div(c1["W"], c1["H"], c1["X"], 0.0, 200, "refcompat")
    assert it == 200
    gold = oracle.kl(c1["Wtest"], c1["Htest"], c1["X"])
    assert abs(cost - gold) / gold < 1e-5
    assert survivors(W) == GOLDEN_SURVIVORS
    assert fro_rel(W, c1["Wtest"]) < 2e-3
    assert fro_rel(H, c1["Htest"]) < 2e-3


def test_exact_known_answer(c1):
    """exact semantics, 100 iterations: KL = 105 982.05, fp32 within 1e-4 of fp64."""
    W, H, it, cost = oracle.update_div(c1["W"], c1["H"], c1["X"], 0.0, 100, "exact")
    assert abs(cost - 105982.05) / 105982.05 < 1e-5
    W64, H64, _, cost64 = oracle.update_div(c1["W"], c1["H"], c1["X"], 0.0, 100, "exact", acc64=True)
    assert abs(cost - cost64) / cost64 < 1e-5
    assert max_rel(W, W64) < 1e-4 and max_rel(H, H64) < 1e-4


def test_exact_does_not_match_golden(c1):
    """the golden encodes the reference's defects: the textbook update must NOT reproduce it"""
    W, H, _, _ = oracle.update_div(c1["W"], c1["H"], c1["X"], 0.0, 20, "exact")
    assert len(survivors(W)) == 128


@pytest.mark.parametrize("shape", [(33, 17, 5), (64, 64, 32), (130, 70, 40), (1, 9, 1)])
def test_kl_monotone_and_ragged_shapes(shape):
    """Paper eq.(2): the multiplicative update never increases the KL divergence."""
    M, N, K = shape
    W, H, X = rand_problem(M, N, K, seed=7)
    prev = oracle.kl(W, H, X)
    for _ in range(5):
        W, H, it, cost = oracle.update_div(W, H, X, 0.0, 1, "exact")
        assert it == 1 and np.isfinite(cost)
        assert cost <= prev * (1 + 1e-6)
        prev = cost
    assert (W >= 0).all() and (H >= 0).all()


def test_half_steps_compose():
    W, H, X = rand_problem(96, 80, 24, seed=3)
    H1 = oracle.h_step(W, H, X)
    W1 = oracle.w_step(W, H1, X)
    W2, H2, _, _ = oracle.update_div(W, H, X, 0.0, 1, "exact")
    assert np.array_equal(W1, W2) and np.array_equal(H1, H2)


def test_quotient_matches_numpy():
    W, H, X = rand_problem(50, 40, 12, seed=5)
    Y = oracle.quotient(W, H, X, acc64=True)
    ref = X.astype(np.float64) / np.maximum(W.astype(np.float64) @ H.astype(np.float64), 2.2204e-16)
    assert max_rel(Y, ref) < 1e-6


def test_zero_inputs_are_clamped_to_eps():
    W, H, X = rand_problem(40, 30, 8, seed=11)
    X[::3, ::2] = 0.0; W[5, :] = 0.0; H[:, 4] = 0.0
    W1, H1, _, cost = oracle.update_div(W, H, X, 0.0, 3, "exact")
    assert np.isfinite(W1).all() and np.isfinite(H1).all() and np.isfinite(cost)


def test_convergence_threshold_stops_early():
    rng = np.random.RandomState(2)
    Wt = rng.rand(60, 4).astype(np.float32); Ht = rng.rand(4, 50).astype(np.float32)
    X = np.asfortranarray(Wt @ Ht)
    W, H, _ = rand_problem(60, 50, 4, seed=9)
    _, _, it, _ = oracle.update_div(W, H, X, 0.5, 400, "exact", check_every=25)
    assert it < 400 and it % 25 == 0
    _, _, it0, _ = oracle.update_div(W, H, X, 0.0, 60, "exact")
    assert it0 == 60


def test_refcompat_row_divide_rule():
    """row_div only runs when rows_padded <= 1024 (cuda/matrix.cu:212-217)."""
    W, H, X = rand_problem(64, 96, 32, seed=1)       # Mp = 64 -> divide happens
    Wr = oracle.w_step(W, H, X, "refcompat")
    We = oracle.w_step(W, H, X, "exact")
    assert not np.allclose(Wr, We)                   # sums are subsampled (7 of 64)
    ratio = (We / Wr)[:, 0]
    assert np.allclose(ratio, ratio[0], rtol=1e-4)   # but only by a per-component scale
