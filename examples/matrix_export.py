#!/usr/bin/env python
"""Input generator compatible with the reference's matrix_export.py (same file format, same random stream),
for every BASELINE configuration instead of the one hard-coded there (matrix_export.py:4-7).

    matrix_export.py [--config c1|c2|c3|c4|c5] [--rows M --cols N --rank K --seed S] [--gpus g] [--out DIR]

Files: X.bin (M x N), W.bin (M x K), H.bin (K x N): two int32 (rows, cols), then rows*cols float32 — the bytes of a
C-ordered numpy array, which read_matrix loads column-major (cuda/nmf.cu:188-218), exactly as the reference's script
writes them.  Legacy numpy stream np.random.seed(S); rand(M,N), rand(M,K), rand(K,N) in that order, so `--config c1`
reproduces the reference's X.bin/W.bin/H.bin byte for byte.  Large matrices are generated and written in row blocks
(the stream is sequential, so the bytes do not depend on the block size).  c3 is the planted low-rank target of
BASELINE config 3: X = (A B) .* (1 + 0.05 u), A, B, u from the same stream after W and H.
"""
import argparse
import os
import struct

import numpy as np

CONFIGS = {            # name: (M, N, K, seed)  — SURVEY.md section 8(d) "Synthetic inputs"
    "c1": (4096, 350, 128, 0),
    "c2": (513, 20000, 40, 1),
    "c3": (2049, 1000000, 128, 3),
    "c4": (16384, 262144, 256, 4),
    "c5": (8192, 131072, 256, 5),
}


def write_rand(path, rows, cols, block_bytes=256 << 20):
    with open(path, "wb") as f:
        f.write(struct.pack("ii", rows, cols))
        step = max(1, block_bytes // (4 * cols))
        for r0 in range(0, rows, step):
            f.write(np.random.rand(min(step, rows - r0), cols).astype(np.float32).tobytes())
    print(f"wrote {path} with {rows * cols} elements")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="c1", choices=sorted(CONFIGS))
    ap.add_argument("--rows", type=int); ap.add_argument("--cols", type=int)
    ap.add_argument("--rank", type=int); ap.add_argument("--seed", type=int)
    ap.add_argument("--gpus", type=int, default=1, help="c5 only: the weak-scaling sweep has 131072 columns per GPU")
    ap.add_argument("--planted", action="store_true", help="X = (A B) .* (1 + 0.05 u) instead of uniform (default for c3)")
    ap.add_argument("--out", default=".")
    a = ap.parse_args()
    M, N, K, seed = CONFIGS[a.config]
    if a.config == "c5":
        N *= a.gpus; seed += max(0, a.gpus.bit_length() - 1)
    M, N, K = a.rows or M, a.cols or N, a.rank or K
    seed = seed if a.seed is None else a.seed
    os.makedirs(a.out, exist_ok=True)
    np.random.seed(seed)
    if a.planted or a.config == "c3":
        # the reference's order (X first) cannot hold for a planted X of this size: W, H first, then the planted factors
        write_rand(os.path.join(a.out, "W.bin"), M, K)
        write_rand(os.path.join(a.out, "H.bin"), K, N)
        A = np.random.rand(M, K).astype(np.float32)          # payload index = column-major: element (i, j) of the
        B = np.random.rand(K, N).astype(np.float32)          # LOADED matrix is flat[j * rows + i]
        with open(os.path.join(a.out, "X.bin"), "wb") as f:
            f.write(struct.pack("ii", M, N))
            Al = A.reshape(-1).reshape((M, K), order="F")     # as read_matrix would load A's bytes
            Bl = B.reshape(-1).reshape((K, N), order="F")
            step = max(1, (64 << 20) // (4 * M))
            for c0 in range(0, N, step):                      # the loaded X column block, written column-major = file order
                blk = (Al @ Bl[:, c0:c0 + step]) * (1.0 + 0.05 * (2.0 * np.random.rand(M, min(step, N - c0)) - 1.0))
                f.write(np.asfortranarray(blk.astype(np.float32)).tobytes(order="F"))
        print(f"wrote {os.path.join(a.out, 'X.bin')} with {M * N} elements (planted rank {K})")
    else:
        write_rand(os.path.join(a.out, "X.bin"), M, N)
        write_rand(os.path.join(a.out, "W.bin"), M, K)
        write_rand(os.path.join(a.out, "H.bin"), K, N)


if __name__ == "__main__":
    main()
