"""Two-kernel path vs fused half-iterations (NMF_UMMA_FUSE=0 / 1) over a range of problem sizes: device time per
iteration of update_div (graph replay, 50 iterations).  The data behind the default fusion rule in umma_host.inc."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nmf_gpu_b200.api as api

SHAPES = [(4096, 350, 128), (513, 20000, 40), (513, 2000, 40), (300, 700, 64), (1024, 1024, 64), (1024, 1024, 256),
          (1024, 4096, 128), (2048, 2048, 128), (4096, 1024, 64), (512, 8192, 64), (8192, 512, 128), (16384, 350, 256),
          (2048, 8192, 64), (2048, 8192, 256), (4096, 4096, 256), (8192, 8192, 128), (2049, 50000, 128),
          (16384, 4096, 256), (256, 100000, 64), (200, 300, 64), (350, 4096, 128)]
rng = np.random.RandomState(0)
for M, N, K in SHAPES:
    X = np.asfortranarray(rng.rand(M, N).astype(np.float32))
    W = np.asfortranarray(rng.rand(M, K).astype(np.float32))
    H = np.asfortranarray(rng.rand(K, N).astype(np.float32))
    row = []
    for fuse, steps in (("0", "3"), ("1", "1"), ("1", "3")):
        os.environ["NMF_UMMA_FUSE"] = fuse
        os.environ["NMF_UMMA_FUSE_STEPS"] = steps       # bit 0: W step, bit 1: H step
        best = 1e9
        for rep in range(3):
            Wd, Hd = W.copy(order="F"), H.copy(order="F")
            r = api.update_div(Wd, Hd, X, 0.0, 50)
            best = min(best, r.device_ms / 50 * 1e3)
        row.append(({"03": "two-kernel", "11": "fused W", "13": "fused W+H"}[fuse + steps], best, r.gpu_launches // 50))
    print(f"{M}x{N} K={K}: " + "  ".join(f"{f}: {us:7.1f} ({l})" for f, us, l in row), flush=True)
