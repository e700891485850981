"""Tolerance-based replacement for the reference's test_output.sh (md5 equality is specific to the
GPU + cuBLAS build that produced Wtest.bin/Htest.bin).  usage: compare_output.py DIR [REFDIR]
Compares DIR/Wout.bin, DIR/Hout.bin with REFDIR/Wtest.bin, REFDIR/Htest.bin (default REFDIR = DIR)."""
import sys

import numpy as np


def read_bin(p):
    with open(p, "rb") as f:
        r, c = np.frombuffer(f.read(8), dtype=np.int32)
        return np.frombuffer(f.read(), dtype=np.float32).reshape((r, c), order="F")


if __name__ == "__main__":
    d = sys.argv[1]; rd = sys.argv[2] if len(sys.argv) > 2 else d
    ok = True
    for out, ref in (("Wout.bin", "Wtest.bin"), ("Hout.bin", "Htest.bin")):
        a, b = read_bin(f"{d}/{out}"), read_bin(f"{rd}/{ref}")
        fro = np.linalg.norm(a.astype(np.float64) - b) / np.linalg.norm(b)
        print(f"{out} vs {ref}: shape {a.shape}, Frobenius-relative difference {fro:.3e}")
        ok &= a.shape == b.shape and fro <= 2e-3
    print("Result matches the test data within tolerance" if ok else "Result DIFFERS from the test data")
    sys.exit(0 if ok else 1)
