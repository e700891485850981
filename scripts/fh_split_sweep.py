"""Fused H step: device time per iteration for forced split counts of the rows of W (NMF_UMMA_FHSPLIT), the data
behind the split-selection model in umma_host.inc (fused_h_splits_impl)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nmf_gpu_b200.api as api

SHAPES = [(4096, 350, 128), (513, 20000, 40), (513, 2000, 40), (300, 700, 64), (1024, 1024, 64), (1024, 1024, 256),
          (1024, 4096, 128), (4096, 1024, 64), (8192, 512, 128), (16384, 350, 256), (2048, 8192, 64), (4096, 4096, 256),
          (8192, 8192, 128), (2049, 50000, 128), (16384, 4096, 256), (350, 4096, 128), (16384, 16384, 256), (2049, 200000, 128)]
rng = np.random.RandomState(0)
os.environ["NMF_UMMA_FUSE"] = "1"
for M, N, K in SHAPES:
    X = np.asfortranarray(rng.rand(M, N).astype(np.float32))
    W = np.asfortranarray(rng.rand(M, K).astype(np.float32))
    H = np.asfortranarray(rng.rand(K, N).astype(np.float32))
    ctiles = (M + 31) // 32 * 32
    ctiles = (ctiles + 255) // 256
    row = []
    for s in ("", "1", "2", "3", "4", "6", "8", "12", "16", "32", "64"):
        if s and int(s) > ctiles:
            continue
        if s:
            os.environ["NMF_UMMA_FHSPLIT"] = s
        else:
            os.environ.pop("NMF_UMMA_FHSPLIT", None)
        best = 1e9
        for rep in range(3):
            Wd, Hd = W.copy(order="F"), H.copy(order="F")
            r = api.update_div(Wd, Hd, X, 0.0, 30)
            best = min(best, r.device_ms / 30 * 1e3)
        row.append((s or "auto", best))
    print(f"{M}x{N} K={K}: " + "  ".join(f"{s}:{us:7.1f}" for s, us in row), flush=True)
