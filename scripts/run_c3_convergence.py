"""BASELINE config C3 as specified: spectrogram-shaped NMF 2049 x 1,000,000, K = 128 on one B200, planted
low-rank target with 5 % multiplicative noise generated on the device, uniform W0/H0, convergence
threshold 1e-6 on the relative KL decrease checked every 25 iterations, cap 2000 iterations.
    python scripts/run_c3_convergence.py [N]      (N defaults to 1000000)"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nmf_gpu_b200.api as api

M, N, K = 2049, int(sys.argv[1]) if len(sys.argv) > 1 else 1000000, 128
with api.Session(M, N, K) as s:
    s.fill_uniform(3, which=3)              # W0, H0
    s.fill_planted(3, K, 0.05)              # X
    c0 = s.cost()
    t0 = time.perf_counter()
    res = s.run(1e-6, 2000)
    dt = time.perf_counter() - t0
    print(json.dumps({"config": f"C3 {M}x{N} K={K} planted rank {K}, noise 0.05, thresh 1e-6, check every 25, cap 2000",
                      "iterations": res.iterations, "converged": res.converged, "cost_start": c0, "cost_end": res.cost,
                      "device_ms": res.device_ms, "ms_per_iteration": res.device_ms / max(1, res.iterations),
                      "effective_tflops": 8.0 * M * N * K * res.iterations / (res.device_ms * 1e-3) / 1e12,
                      "wall_s": dt}))
