"""Shared error metrics (BASELINE.md 'Parity gates')."""
import numpy as np


def max_rel(a, ref):
    """max |a-ref| / max(|ref|, 1e-6 * max|ref|) — relative error that ignores near-underflow entries."""
    ref = np.asarray(ref, dtype=np.float64); a = np.asarray(a, dtype=np.float64)
    den = np.maximum(np.abs(ref), 1e-6 * np.abs(ref).max())
    return float((np.abs(a - ref) / den).max())


def fro_rel(a, ref):
    ref = np.asarray(ref, dtype=np.float64); a = np.asarray(a, dtype=np.float64)
    return float(np.linalg.norm(a - ref) / np.linalg.norm(ref))


def survivors(W, rel=1e-12):
    """indices of components (columns of W) that have not collapsed: column mass above rel * the largest.
    (The reference's golden holds exact zeros there; its build on B200 leaves denormal-scale residue.)"""
    mass = np.abs(np.asarray(W, dtype=np.float64)).sum(axis=0)
    return sorted(np.nonzero(mass > rel * mass.max())[0].tolist())


def rand_problem(M, N, K, seed):
    rng = np.random.RandomState(seed)
    X = np.asfortranarray(rng.rand(M, N).astype(np.float32))
    W = np.asfortranarray(rng.rand(M, K).astype(np.float32))
    H = np.asfortranarray(rng.rand(K, N).astype(np.float32))
    return W, H, X
