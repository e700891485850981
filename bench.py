#!/usr/bin/env python
"""bench.py — NMF iterations/s and effective TFLOP/s (8*M*N*K flop per iteration) of the update_div
hot path on 1/2/4/8 B200 (BASELINE.json `metric`).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload c4|c3|c5|c5k64|c5k512|c5w|c5wk64|c5wk512|c2|c1|tiny]

A "step" is one NMF iteration (H half-step + W half-step).  Workload at every N: BASELINE config C4
"large synthetic NMF 16384x262144 K=256 column-sharded across 1/2/4/8 B200" — total work fixed, so
`scaling` is "strong".  One process per GPU (torchrun); X and H are sharded by columns, W replicated,
one NCCL allreduce of [numerator | row sums] per iteration.

  value : iterations/s, device timed (cudaEvents around the loop inside the library), inputs
          generated in HBM before the timed region; max over ranks.
  e2e   : the same iterations through the reference-facing call with HOST buffers — update_div(W,H,X)
          at N=1, upload+run+download of the rank's shard at N>1 — H2D/D2H inside the timed region.
  roofline / cpu_baseline / clocks / gpu_launches: see DESIGN.md "Measurement".
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {                     # name: (M, N, K, seed, description)
    "c4": (16384, 262144, 256, 4, "C4 large synthetic NMF 16384x262144 K=256"),
    "c5": (8192, 131072, 256, 5, "C5 8192x131072 K=256"),
    "c3": (2049, 1000000, 128, 3, "C3 spectrogram NMF 2049x1000000 K=128"),
    "c3s": (2049, 262144, 128, 3, "C3-shape 2049x262144 K=128"),
    "c5k64": (8192, 131072, 64, 5, "C5 8192x131072 K=64"),
    "c5k512": (8192, 131072, 512, 5, "C5 8192x131072 K=512"),
    # BASELINE config 5 as a WEAK-scaling sweep: 131072 columns PER GPU (N = 131072 * g), seed 5 + log2(g)
    "c5w": (8192, 131072, 256, 5, "C5 weak 8192x(131072*g) K=256"),
    "c5wk64": (8192, 131072, 64, 5, "C5 weak 8192x(131072*g) K=64"),
    "c5wk512": (8192, 131072, 512, 5, "C5 weak 8192x(131072*g) K=512"),
    "c2": (513, 20000, 40, 1, "C2 audio spectrogram 513x20000 K=40"),
    "c1": (4096, 350, 128, 0, "C1 bundled shape 4096x350 K=128"),
    "mid": (8192, 32768, 256, 7, "mid 8192x32768 K=256 (profiling size)"),
    "tiny": (1024, 4096, 64, 9, "tiny 1024x4096 K=64"),
}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(bf16=d["bf16_tflops"], bf16_sustained=d.get("bf16_tflops_sustained", d["bf16_tflops"]),
                    hbm=d["hbm_gbs"], source="measured (MEASURED_PEAKS.json)")
    return dict(bf16=1590.0, bf16_sustained=1400.0, hbm=6650.0, source="fallback (B200_PROFILING.md)")


class ClockSampler:
    """nvidia-smi SM clocks + throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True); self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate(); self.t.join(timeout=2)
        sm = [float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 9 for n, v in zip(names, r[5:9]) if v.lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def shard(N, rank, world):
    from nmf_gpu_b200.sharding import column_shard
    return column_shard(N, rank, world)


def cpu_baseline(M, N, K, seed, target_s=12.0):
    """The oracle (C++/OpenMP port of the iteration; the reference has no CPU path) on this box's
    host cores, on a bounded column sample of the same workload, scaled to whole-workload iterations/s."""
    from oracle import oracle
    ns = int(min(N, max(256, 2048 * (16384 * 256) // max(1, M * K))))
    rng = np.random.RandomState(seed)
    X = np.asfortranarray(rng.rand(M, ns).astype(np.float32))
    W = np.asfortranarray(rng.rand(M, K).astype(np.float32))
    H = np.asfortranarray(rng.rand(K, ns).astype(np.float32))
    s1, th = oracle.time_iterations(W, H, X, 1)
    iters = int(max(1, min(50, target_s / max(s1, 1e-3))))
    s, th = oracle.time_iterations(W, H, X, iters)
    per_iter_full = (s / iters) * (N / ns)
    return {"value": 1.0 / per_iter_full, "unit": "iterations/s", "cores": th, "kind": "port",
            "sample": f"{iters} iterations on the first {ns} of {N} columns (M={M}, K={K}), scaled by {N}/{ns}",
            "tflops": 8.0 * M * N * K / per_iter_full / 1e12}


def run_reference_arm(args, M, N, K, seed, desc, rank, world):
    """--impl reference: the path's CPU implementation on the host cores.  The reference itself has no
    CPU path (every op is a CUDA kernel / cuBLAS call), so this is the oracle port; rank 0 only."""
    if rank != 0:
        return
    t0 = time.time()
    cb = cpu_baseline(M, N, K, seed, target_s=max(4.0, 2.0 * (args.steps + args.warmup)))
    line = {"impl": "reference", "metric": "nmf_iterations_per_s", "value": cb["value"], "unit": "iterations/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 / cb["value"],
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": desc, "M": M, "N": N, "K": K},
            "effective_tflops": cb["tflops"], "cpu_baseline": cb,
            "e2e": {"value": cb["value"], "unit": "iterations/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "wall_s": time.time() - t0}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--workload", default=os.environ.get("NMF_BENCH_WORKLOAD", "c4"))
    ap.add_argument("--engine", default="tcgen05")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    M, N, K, seed, desc = WORKLOADS[args.workload]
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    weak = args.workload.startswith("c5w")
    if weak:                                                  # per-GPU work fixed: the column count grows with the GPU count
        N *= world; seed += max(0, world.bit_length() - 1)
        desc = desc.replace("(131072*g)", str(N)) + f" (g={world})"
    if args.impl == "reference":
        return run_reference_arm(args, M, N, K, seed, desc, rank, world)
    args.warmup = max(args.warmup, 3)

    import torch
    import nmf_gpu_b200.api as api
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        if dist is None:
            return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    n_off, n_loc = shard(N, rank, world)
    opt = api.options(engine=args.engine, device=local)
    free0 = torch.cuda.mem_get_info()[0]
    sess = api.Session(M, n_loc, K, n_offset=n_off, n_global=N, opt=opt)
    hbm_resident = int(free0 - torch.cuda.mem_get_info()[0])  # the session's device buffers (X, W, H + pairs, partials; no quotient when fused)

    def join(session):
        """fresh ncclUniqueId from rank 0, broadcast over torch.distributed, then ncclCommInitRank"""
        if world == 1:
            return
        uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            uid = torch.frombuffer(bytearray(api.nccl_unique_id()), dtype=torch.uint8).cuda()
        dist.broadcast(uid, 0)
        session.join(rank, world, uid.cpu().numpy().tobytes())

    join(sess)
    sess.fill_uniform(seed)                                   # inputs resident in HBM before timing

    # ---- device-timed region: W warm-up steps, then exactly K steps -------------------------
    sess.run(0.0, args.warmup)
    sampler = ClockSampler(local);
    if rank == 0:
        sampler.start()
    barrier()
    res = sess.run(0.0, args.steps)
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    ms = max_over_ranks(res.device_ms)
    ms_per_step = ms / args.steps
    value = 1e3 / ms_per_step
    flop_iter = 8.0 * M * N * K
    launches = res.gpu_launches

    # ---- per-kernel timing for the roofline (eager run with cudaEvent pairs per phase) --------
    t = np.zeros(10)
    barrier()
    sess.run(0.0, 3, t=t)
    barrier()
    fused = sess.fused_steps()                                # bit 0: W step, bit 1: H step run as one kernel each
    nq = 2 - bin(fused).count("1")                            # stand-alone quotient launches per iteration
    quot_ms = t[0] * 1e3 / (nq * 3) if nq else 0.0
    gemm2_ms = t[3] * 1e3 / (2 * 3)                           # the H-step and W-step GEMM kernels (fused: incl. their quotient)
    pk = peaks()
    peak3 = pk["bf16_sustained"] / 2.0 / 3.0                  # TF32 dense = bf16/2; 3 MMAs per product
    traffic = None
    try:   # DRAM bytes per launch of the same kernel from the committed ncu --set full capture (profiles/)
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        traffic = tj.get(args.workload + ("+fused" if fused == 3 else ""), {}).get(str(world))
    except (OSError, ValueError):
        pass
    if fused == 3:
        # dominant kernel = k_fused (quotient GEMM + consuming GEMM in one launch): 4*M*n_loc*K algorithmic flop per launch
        ach = 4.0 * M * n_loc * K / (gemm2_ms * 1e-3) / 1e12 if gemm2_ms > 0 else 0.0
        kernel = "k_fused<W step | H step> (W*H GEMM -> X./max(.,EPS) -> TF32 pair in smem -> second GEMM, 3xTF32)"
        launch_ms = gemm2_ms
        alg_bytes = 4.0 * M * n_loc + 8.0 * K * (M + n_loc) + 12.0 * K * min(M, n_loc)   # X read once; W, H pairs; the small output
    else:
        # dominant kernel = k_umma<QUOT> (W*H with the fused clamp/divide), 2*M*n_loc*K algorithmic flop per launch
        ach = 2.0 * M * n_loc * K / (quot_ms * 1e-3) / 1e12 if quot_ms > 0 else 0.0
        kernel = "k_umma<QUOT> (W*H GEMM + fused X./max(.,EPS), 3xTF32)"
        launch_ms = quot_ms
        alg_bytes = 8.0 * M * n_loc + 8.0 * K * (M + n_loc)   # X read + Y written once, W and H (hi, lo) read
    # The denominator is MEASURED_PEAKS' cuBLAS bf16 figure / 2 (TF32) / 3 (three MMAs per product): the sustained one
    # for a kernel inside a long step.  The fused kernels draw less power per flop than that cuBLAS run (no M x N
    # round trip through HBM), so they clock higher and pass it; the burst figure is the ceiling then.  Both are
    # power-capped clocks, so `frac_at_clock` also gives achieved / (148 SMs x 4096 TF32 flop/clk x the SM clock sampled
    # during the timed region / 3), i.e. the tensor-pipe utilisation ncu reports.
    peak_kind, peak_src = "sustained", pk["bf16_sustained"]
    if ach > peak3:
        peak3, peak_kind, peak_src = pk["bf16"] / 2.0 / 3.0, "burst", pk["bf16"]
    nominal3 = 148 * 4096 * float((clocks or {}).get("sm_max_mhz") or 1965.0) * 1e6 / 3.0 / 1e12
    if ach > peak3:    # short runs (N = 8: < 100 ms) finish before the power cap bites: the ceiling is the tensor pipe at the maximum clock
        peak3, peak_kind, peak_src = nominal3, "nominal (148 SMs x 4096 TF32 flop/clk x max SM clock)", nominal3 * 6.0
    clk_mhz = (clocks or {}).get("sm_mhz") if rank == 0 else None
    frac_at_clock = ach / (148 * 4096 * clk_mhz * 1e6 / 3.0 / 1e12) if clk_mhz else None
    roofline = {"bound": "tensor", "kernel": kernel,
                "achieved": ach, "peak": peak3, "unit": "TFLOP/s", "frac": ach / peak3, "traffic": traffic,
                "algorithmic_hbm_bytes": alg_bytes, "frac_at_clock": frac_at_clock,
                "peak_note": f"{pk['source']}: bf16 {peak_kind} {peak_src:.0f} TF/s / 2 (TF32) / 3 (3xTF32 passes)",
                "launch_ms": launch_ms, "other_gemm_launch_ms": gemm2_ms if fused != 3 else None, "fused_steps": fused,
                "whole_iteration_frac": (flop_iter / world / (ms_per_step * 1e-3) / 1e12) / peak3,   # same denominator
                "phase_seconds_3_iterations": [float(v) for v in t]}
    if world == 1:
        sess.close()      # N > 1: kept until the e2e run has shared its communicator (a process creates NCCL once)

    # ---- e2e: host buffers, H2D and D2H inside the timed region --------------------------------
    e2e = None
    if not args.no_e2e:
        g = torch.Generator(); g.manual_seed(seed * 1000 + rank)
        Xh = torch.empty((n_loc, M), dtype=torch.float32, pin_memory=True).uniform_(0, 1, generator=g)
        Wh = torch.empty((K, M), dtype=torch.float32, pin_memory=True).uniform_(0, 1, generator=torch.Generator().manual_seed(seed))
        Hh = torch.empty((n_loc, K), dtype=torch.float32, pin_memory=True).uniform_(0, 1, generator=g)
        Xn, Wn, Hn = Xh.numpy().T, Wh.numpy().T, Hh.numpy().T    # column-major views (M x n), (M x K), (K x n)
        barrier()
        t0 = time.perf_counter()
        if world == 1:
            r2 = api.update_div(Wn, Hn, Xn, 0.0, args.steps, opt=opt)     # the reference-facing call
        else:
            s2 = api.Session(M, n_loc, K, n_offset=n_off, n_global=N, opt=opt)
            s2.share_comm(sess)                                            # the process-wide communicator
            s2.upload(Wn, Hn, Xn); r2 = s2.run(0.0, args.steps); s2.download(); s2.close()
        torch.cuda.synchronize()
        dt = max_over_ranks(time.perf_counter() - t0)
        barrier()
        e2e = {"value": args.steps / dt, "unit": "iterations/s",
               "h2d_bytes_per_step": int(4 * (M * n_loc + M * K + K * n_loc) / args.steps),
               "d2h_bytes_per_step": int(4 * (M * K + K * n_loc) / args.steps),
               "seconds": dt, "call": "update_div(W,H,X,0,steps)" if world == 1 else "Session.upload/run/download per rank",
               "loop_device_ms": r2.device_ms}
        del Xh, Wh, Hh
    if world > 1:
        sess.close()

    if rank == 0:
        line = {"metric": "nmf_iterations_per_s", "value": value, "unit": "iterations/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
                "scaling": "weak" if weak else "strong", "vs_baseline": None, "dtype": "f32 (3xTF32 tensor-core products, fp32 accumulate)",
                "data": "synthetic",
                "config": {"workload": desc, "M": M, "N": N, "K": K, "columns_per_gpu": n_loc,
                           "parallelism": f"column-sharded x{world}", "engine": args.engine,
                           "hbm_resident_bytes_per_gpu": hbm_resident,
                           "l2": "inputs (X alone >= 2 GB per GPU) far exceed the 126 MB L2; no flush needed"},
                "effective_tflops": flop_iter / (ms_per_step * 1e-3) / 1e12,
                "gpu_launches": int(launches), "e2e": e2e, "roofline": roofline, "clocks": clocks}
        if world == 1 and not args.no_cpu:
            line["cpu_baseline"] = cpu_baseline(M, N, K, seed)
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
