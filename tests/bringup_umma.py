"""Bring-up harness for the tcgen05 GEMM kernel (not a pytest file).

Runs every (geometry, passes, descriptor variant, shape) case of nmf_dbg_gemm in its OWN subprocess
under a timeout: a wrong descriptor can trap and poison the CUDA context, and that must not take
the remaining cases (or the GPU box) down.  Prints one line per case.

    python tests/bringup_umma.py            # all cases
    python tests/bringup_umma.py case 0 3 0 128 256 64     # one case (used internally)
"""
import json
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def one_case(which, passes, variant, M, N, K):
    os.environ["NMF_UMMA_VARIANT"] = str(variant)
    import nmf_gpu_b200.api as api
    rng = np.random.RandomState(1000 * which + M + N + K)
    shapes = {0: ((M, K), (K, N)), 1: ((M, K), (M, N)), 2: ((M, N), (K, N))}[which]
    A = np.asfortranarray(rng.rand(*shapes[0]).astype(np.float32) - 0.3)
    B = np.asfortranarray(rng.rand(*shapes[1]).astype(np.float32) - 0.3)
    C = api.dbg_gemm(which, A, B, M, N, K, passes)
    A64, B64 = A.astype(np.float64), B.astype(np.float64)
    ref = {0: lambda: A64 @ B64, 1: lambda: A64.T @ B64, 2: lambda: A64 @ B64.T}[which]()
    err = np.abs(C - ref) / np.abs(ref).max()
    out = dict(which=which, passes=passes, variant=variant, M=M, N=N, K=K, max_rel=float(np.nanmax(err)),
               nan=int(np.isnan(C).sum()), frac_ok=float((err < 1e-3).mean()))
    if out["max_rel"] > 1e-3:
        bad = np.argwhere(~(err < 1e-3))
        out["first_bad"] = bad[:4].tolist()
        out["bad_rows"] = int(len(set(bad[:, 0].tolist()))); out["bad_cols"] = int(len(set(bad[:, 1].tolist())))
    print("CASE " + json.dumps(out))


def acc_probe(R):
    """How does the tensor core round its fp32 accumulator?  Positive inputs with 6-bit mantissas:
    every product is exact in fp32, the split's lo parts are zero, so any error is accumulation
    rounding.  Mean signed error ~0 => round-to-nearest; negative ~ -n/2 ulp => truncation."""
    import nmf_gpu_b200.api as api
    rng = np.random.RandomState(5)
    q = lambda a: (np.floor(a * 64) / 64 + 1.0 / 64).astype(np.float32)       # in (0,1], 6 bits
    K, N = 128, 256
    A = np.asfortranarray(q(rng.rand(R, K))); B = np.asfortranarray(q(rng.rand(R, N)))
    out = dict(probe="acc", R=R)
    for passes in (1, 3):
        C = api.dbg_gemm(1, A, B, R, N, K, passes)
        ref = A.astype(np.float64).T @ B.astype(np.float64)
        rel = (C.astype(np.float64) - ref) / ref
        out[f"p{passes}_mean_signed"] = float(rel.mean()); out[f"p{passes}_max_abs"] = float(np.abs(rel).max())
    # fp32 round-to-nearest sequential reference for scale
    out["ulp"] = 2.0 ** -24
    print("CASE " + json.dumps(out))


def tf32_rounding_probe():
    """Does tcgen05.mma.kind::tf32 truncate or round the 13 dropped mantissa bits? (1-pass, raw inputs
    are pre-masked by the split kernel, so feed values whose lo part is exactly half an ulp...)"""
    # with the explicit hi/lo split the hardware only ever sees hi (exact) and lo (<= 13 significant
    # bits, re-truncated to 11): nothing to probe for correctness; kept as a note.
    return None


CASES = [
    # which, passes, variant, M, N, K
    (0, 1, 0, 128, 256, 32), (0, 1, 1, 128, 256, 32),
    (1, 1, 0, 128, 256, 128),
    (2, 1, 0, 128, 64, 256), (2, 1, 1, 128, 64, 256),
    (0, 3, 0, 128, 256, 64), (1, 3, 0, 256, 256, 128), (2, 3, 0, 256, 128, 256),
    (0, 3, 0, 160, 100, 96), (1, 3, 0, 160, 100, 96), (2, 3, 0, 160, 100, 96),
    (0, 3, 0, 512, 1024, 256), (1, 3, 0, 512, 1024, 256), (2, 3, 0, 512, 1024, 256),
]

if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "case":
        one_case(*[int(v) for v in sys.argv[2:8]])
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "acc":
        acc_probe(int(sys.argv[2]))
        sys.exit(0)
    for c in CASES + [("acc", 512), ("acc", 4096), ("acc", 16384)]:
        cmd = [sys.executable, os.path.abspath(__file__)] + (["case"] if c[0] != "acc" else []) + [str(v) for v in c]
        try:
            r = subprocess.run(cmd, capture_output=True, text=True, timeout=120)
            lines = [l for l in r.stdout.splitlines() if l.startswith("CASE ")]
            if lines:
                print(lines[-1], flush=True)
            else:
                print("FAIL", c, "rc", r.returncode, (r.stderr or r.stdout)[-400:].replace("\n", " | "), flush=True)
        except subprocess.TimeoutExpired:
            print("TIMEOUT", c, flush=True)
