"""Builds tests/golden/golden_c1.npz from the reference's own golden vectors.

Run in the build container (needs /root/reference):  python tests/golden/make_golden.py

The reference ships Wtest.bin / Htest.bin (outputs of its CUDA program after 200 iterations on
the inputs of matrix_export.py) and compares them by md5 (test_output.sh).  They are data, not
source; they are stored here compressed (87 % of the entries are exact zeros) so the GPU box,
which has no /root/reference, can check against them.

File format (README.md:31-36, cuda/nmf.cu:188-259): int32 rows, int32 cols, float32 payload
interpreted COLUMN-MAJOR.
"""
import os
import sys

import numpy as np

REF = os.environ.get("NMF_REFERENCE", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))


def read_bin(path):
    with open(path, "rb") as f:
        r, c = np.frombuffer(f.read(8), dtype=np.int32)
        a = np.frombuffer(f.read(), dtype=np.float32)
    assert a.size == r * c, (path, r, c, a.size)
    return a.reshape((r, c), order="F").copy(order="F")


def c1_inputs():
    """matrix_export.py:4-7 verbatim (legacy numpy seed 0), read back the way read_matrix does:
    the script writes C-order bytes under a (rows, cols) header and the reference reads the
    payload column-major, so the loaded matrix is flat.reshape((r, c), order='F')."""
    np.random.seed(0)
    X = np.random.rand(4096, 350).astype(np.float32)
    W = np.random.rand(4096, 128).astype(np.float32)
    H = np.random.rand(128, 350).astype(np.float32)
    as_loaded = lambda a: a.reshape(-1).reshape(a.shape, order="F").copy(order="F")
    return as_loaded(W), as_loaded(H), as_loaded(X)


if __name__ == "__main__":
    Wt = read_bin(os.path.join(REF, "Wtest.bin"))
    Ht = read_bin(os.path.join(REF, "Htest.bin"))
    out = os.path.join(HERE, "golden_c1.npz")
    np.savez_compressed(out, Wtest=Wt, Htest=Ht)
    print("wrote", out, Wt.shape, Ht.shape, os.path.getsize(out), "bytes")
    sys.exit(0)
