"""One-sided (truncation) error of the chunked tcgen05 reductions vs the promotion period.
Run on the GPU box:  python scripts/bias_probe.py"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nmf_gpu_b200.api as api

rng = np.random.RandomState(3)
R = 16384
A1 = np.asfortranarray(rng.rand(R, 64).astype(np.float32)); B1 = np.asfortranarray(rng.rand(R, 96).astype(np.float32))
ref1 = A1.astype(np.float64).T @ B1.astype(np.float64)
A2 = np.asfortranarray(rng.rand(256, R).astype(np.float32)); B2 = np.asfortranarray(rng.rand(64, R).astype(np.float32))
ref2 = A2.astype(np.float64) @ B2.astype(np.float64).T
A0 = np.asfortranarray(rng.rand(256, 256).astype(np.float32)); B0 = np.asfortranarray(rng.rand(256, 96).astype(np.float32))
ref0 = A0.astype(np.float64) @ B0.astype(np.float64)
# fp32 reference: plain fp32 accumulation error level for comparison
f32 = (A1.T @ B1).astype(np.float64)
print("numpy fp32 matmul      : mean rel %+.2e  max |rel| %.2e" % (((f32 - ref1) / ref1).mean(), np.abs((f32 - ref1) / ref1).max()))
for chunk in (1, 2, 4, 8, 16):
    os.environ["NMF_UMMA_CHUNK"] = str(chunk)
    C1 = api.dbg_gemm(1, A1, B1, R, 96, 64, passes=3); e1 = (C1 - ref1) / ref1
    C2 = api.dbg_gemm(2, A2, B2, 256, R, 64, passes=3); e2 = (C2 - ref2) / ref2
    print("chunk %2d  G2-type (R=%d): mean rel %+.2e max %.2e | G3-type: mean rel %+.2e max %.2e"
          % (chunk, R, e1.mean(), np.abs(e1).max(), e2.mean(), np.abs(e2).max()))
C0 = api.dbg_gemm(0, A0, B0, 256, 96, 256, passes=3); e0 = (C0 - ref0) / ref0
print("G1-type whole-tile K=256: mean rel %+.2e max %.2e" % (e0.mean(), np.abs(e0).max()))
