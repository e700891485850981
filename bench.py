#!/usr/bin/env python
"""bench.py — NMF iterations/s and effective TFLOP/s (8*M*N*K flop per iteration) of the update_div
hot path on 1/2/4/8 B200 (BASELINE.json `metric`).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
                  [--workload c4|c3|c5|c5k64|c5k512|c5w|c5wk64|c5wk512|c2|c2r4|c2r8|c1|mid|tiny]

A "step" is one NMF iteration (H half-step + W half-step).  Workload at every N: BASELINE config C4
"large synthetic NMF 16384x262144 K=256 column-sharded across 1/2/4/8 B200" — total work fixed, so
`scaling` is "strong".  One process per GPU (torchrun); X and H are sharded by columns, W replicated,
one NCCL allreduce of [numerator | row sums] per iteration.

  value : iterations/s, device timed (cudaEvents around the loop inside the library), inputs
          generated in HBM before the timed region; max over ranks.
  e2e   : the same iterations through the reference-facing call with HOST buffers — update_div(W,H,X)
          at N=1, upload+run+download of the rank's shard at N>1 — H2D/D2H inside the timed region.
  roofline       : the dominant kernel against ONE fixed denominator — MEASURED_PEAKS.json's burst bf16 figure / 2 (TF32)
                   / 3 (3xTF32 passes) — plus frac_sustained / frac_nominal / frac_at_clock as separate keys.
  sustained      : the same loop re-run for >= 2 s of device time (power-capped steady state), clocks sampled.
  reference_cuda : the reference's own CUDA program (baseline/_ref/nmf_ref_patched: cuBLAS SGEMM x4 + its SIMT kernels,
                   patched only so that it can run this size) timed on the same GPU right after ours (N = 1).
  parity_vs_1gpu : N > 1: a small problem run sharded over the N ranks and unsharded on rank 0, compared.
  cpu_baseline / clocks / gpu_launches: see DESIGN.md "Measurement".
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

WORKLOADS = {                     # name: (M, N, K, seed, description[, restarts])
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
    # batched random restarts: R initialisations of C2 in the same launches; value counts iteration x restart
    "c2r4": (513, 20000, 40, 1, "C2 audio spectrogram 513x20000 K=40, 4 batched restarts", 4),
    "c2r8": (513, 20000, 40, 1, "C2 audio spectrogram 513x20000 K=40, 8 batched restarts", 8),
    "c1": (4096, 350, 128, 0, "C1 bundled shape 4096x350 K=128"),
    "c4s8": (16384, 32768, 256, 4, "one rank's shard of C4 at 8 GPUs: 16384x32768 K=256 (profiling size)"),
    "mid": (8192, 32768, 256, 7, "mid 8192x32768 K=256 (profiling size)"),
    "tiny": (1024, 4096, 64, 9, "tiny 1024x4096 K=64"),
}
L2_BYTES = 126e6


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(bf16=d["bf16_tflops"], bf16_sustained=d.get("bf16_tflops_sustained", d["bf16_tflops"]),
                    hbm=d["hbm_gbs"], source="MEASURED_PEAKS.json")
    return dict(bf16=1590.0, bf16_sustained=1400.0, hbm=6650.0, source="fallback of B200_PROFILING.md (MEASURED_PEAKS.json absent)")


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
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate(); self.t.join(timeout=2)
        sm = [float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        pw = [float(r[3]) for r in self.rows if len(r) >= 9 and r[3].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 9 for n, v in zip(names, r[5:9]) if v.lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm), "power_w_max": max(pw) if pw else None}


def shard(N, rank, world):
    from nmf_gpu_b200.sharding import column_shard
    return column_shard(N, rank, world)


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def bind_near_gpu(index):
    """Run this process (and first-touch its pinned host buffers) on the CPUs next to GPU `index` (NVML's ideal affinity):
    torchrun does not place its workers, and a shard uploaded from the far socket crosses the inter-socket link
    (round 1: 11 GB/s per rank at N = 4).  Returns the previous affinity (restored before the CPU baseline runs)."""
    try:
        import pynvml
        prev = os.sched_getaffinity(0)
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = {64 * w + b for w, m in enumerate(mask) for b in range(64) if (m >> b) & 1} & prev
        if cpus:
            os.sched_setaffinity(0, cpus)
        return prev, sorted(cpus)
    except Exception:
        return None, None


def cpu_baseline(M, N, K, seed, target_s=12.0):
    """The oracle (C++/OpenMP port of the iteration; the reference has no CPU path) on this box's
    host cores, on a bounded column sample of the same workload, scaled to whole-workload iterations/s.
    All host cores, also under torchrun (which exports OMP_NUM_THREADS=1 to its workers)."""
    from oracle import oracle
    oracle.set_threads(host_threads())
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
            "scaled": ns < N, "sample_columns": ns, "sample_iterations": iters, "sample_seconds": s,
            "sample": f"{iters} iterations on the first {ns} of {N} columns (M={M}, K={K}), time scaled by {N}/{ns}",
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
            "effective_tflops": cb["tflops"], "cpu_baseline": cb, "scaled": cb["scaled"], "sample_columns": cb["sample_columns"],
            "e2e": {"value": cb["value"], "unit": "iterations/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "wall_s": time.time() - t0}
    print(json.dumps(line), flush=True)


def reference_cuda(M, N, K, steps, warmup, device):
    """The reference's own CUDA program on this GPU for this size (baseline/_ref/nmf_ref_patched, built by
    baseline/make_ref_patched.py from /root/reference/cuda: cuBLAS SGEMM x4 + 12 SIMT passes per iteration, exact sums).
    Inputs generated on the device; cudaEvent time of its graph replay loop."""
    exe = os.path.join(ROOT, "baseline", "_ref", "nmf_ref_patched")
    if not os.path.exists(exe):
        return {"unavailable": "baseline/_ref/nmf_ref_patched not built (python baseline/make_ref_patched.py needs /root/reference)"}
    need = 4.0 * (2.0 * M * N + 2.0 * K * N + 2.0 * M * K)          # X, Z + H, WtZ + W, ZHt
    try:
        import torch
        if need > 0.9 * torch.cuda.mem_get_info()[0]:
            return {"unavailable": f"needs {need / 1e9:.1f} GB of HBM"}
    except Exception:
        pass
    vis = [v for v in os.environ.get("CUDA_VISIBLE_DEVICES", "").split(",") if v]
    env = dict(os.environ, NVIDIA_TF32_OVERRIDE="0", CUDA_VISIBLE_DEVICES=vis[device] if device < len(vis) else str(device))
    env["LD_LIBRARY_PATH"] = "/usr/local/cuda/lib64:" + env.get("LD_LIBRARY_PATH", "")
    try:
        r = subprocess.run([exe, str(M), str(N), str(K), str(steps), "-", str(warmup)], env=env, capture_output=True, text=True, timeout=900)
    except subprocess.TimeoutExpired:
        return {"unavailable": "timed out"}
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    if r.returncode != 0 or not lines:
        return {"unavailable": f"exit {r.returncode}: {(r.stderr or r.stdout)[-300:]}"}
    d = json.loads(lines[-1])
    return {"value": d["iterations_per_s"], "unit": "iterations/s", "ms_per_step": d["ms_per_iteration"], "steps": steps, "warmup": warmup,
            "effective_tflops": 8.0 * M * N * K / (d["ms_per_iteration"] * 1e-3) / 1e12,
            "what": "reference CUDA program (cublasSgemm x4 + its SIMT kernels per iteration, fp32, NVIDIA_TF32_OVERRIDE=0), "
                    "patched for 64-bit sizes / large grids / true sums only (baseline/make_ref_patched.py), same GPU, device-generated inputs"}


def parity_vs_1gpu(api, torch, dist, rank, world, local):
    """30 iterations of a 1024 x (1024*world + 300) x 128 problem: sharded over the ranks vs unsharded on rank 0."""
    from nmf_gpu_b200.sharding import column_shard
    M, N, K, iters = 1024, 1024 * world + 300, 128, 30
    rng = np.random.RandomState(11)
    X = np.asfortranarray(rng.rand(M, N).astype(np.float32)); W = np.asfortranarray(rng.rand(M, K).astype(np.float32))
    H = np.asfortranarray(rng.rand(K, N).astype(np.float32))
    lo, n = column_shard(N, rank, world)
    opt = api.options(device=local)
    s = api.Session(M, n, K, n_offset=lo, n_global=N, opt=opt)
    uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
    if rank == 0:
        uid = torch.frombuffer(bytearray(api.nccl_unique_id()), dtype=torch.uint8).cuda()
    dist.broadcast(uid, 0)
    s.join(rank, world, uid.cpu().numpy().tobytes())
    s.upload(W, H[:, lo:lo + n], X[:, lo:lo + n])
    res = s.run(1e-9, iters)
    Wd, Hd = s.download(); s.close()
    Hfull = torch.zeros((N, K), dtype=torch.float32, device="cuda")
    Hfull[lo:lo + n] = torch.from_numpy(np.ascontiguousarray(Hd.T)).cuda()
    dist.all_reduce(Hfull)
    if rank != 0:
        return None
    Hm = Hfull.cpu().numpy().T
    with api.Session(M, N, K, opt=opt) as s1:
        s1.upload(W, H, X); r1 = s1.run(1e-9, iters); W1, H1 = s1.download()
    rel = lambda a, r: float((np.abs(a.astype(np.float64) - r) / np.maximum(np.abs(r), 1e-6 * np.abs(r).max())).max())
    out = {"problem": f"{M}x{N} K={K}, {iters} iterations, thresh 1e-9", "W_max_rel": rel(Wd, W1), "H_max_rel": rel(Hm, H1),
           "cost_rel": abs(res.cost - r1.cost) / abs(r1.cost), "tolerance": 2e-5}
    out["ok"] = bool(out["W_max_rel"] < 2e-5 and out["H_max_rel"] < 2e-5 and out["cost_rel"] < 1e-5)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--workload", default=os.environ.get("NMF_BENCH_WORKLOAD", "c4"))
    ap.add_argument("--engine", default="tcgen05")
    ap.add_argument("--tf32-passes", type=int, default=3, choices=(1, 3),
                    help="3 = 3xTF32, fp32-equivalent (default, the north star's arithmetic); 1 = single-pass TF32 (~1e-3 accuracy, SURVEY 8(f4)): "
                         "reported with its own dtype string, never as the headline")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-refcuda", action="store_true")
    ap.add_argument("--no-sustained", action="store_true")
    args = ap.parse_args()
    wl = WORKLOADS[args.workload]
    M, N, K, seed, desc = wl[:5]
    R = wl[5] if len(wl) > 5 else 1
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    weak = args.workload.startswith("c5w")
    if weak:                                                  # per-GPU work fixed: the column count grows with the GPU count
        N *= world; seed += max(0, world.bit_length() - 1)
        desc = desc.replace("(131072*g)", str(N)) + f" (g={world})"
    if args.impl == "reference":
        return run_reference_arm(args, M, N, K, seed, desc, rank, world)
    args.warmup = max(args.warmup, 3)

    vis = [v for v in os.environ.get("CUDA_VISIBLE_DEVICES", "").split(",") if v]
    prev_affinity, near_cpus = bind_near_gpu(int(vis[local]) if local < len(vis) and vis[local].isdigit() else local)
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
    opt = api.options(engine=args.engine, device=local, restarts=R, tf32_passes=args.tf32_passes)
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
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    barrier()
    res = sess.run(0.0, args.steps)
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    ms = max_over_ranks(res.device_ms)
    ms_per_step = ms / args.steps
    value = R * 1e3 / ms_per_step                             # restarts: iteration x restart per second
    flop_iter = 8.0 * M * N * K * R
    launches = res.gpu_launches

    # ---- sustained: the same loop for >= 2 s of device time (the power cap engages after ~1 s) ----
    sustained = None
    if not args.no_sustained:
        n_sus = int(min(20000, max(args.steps, np.ceil(2000.0 / ms_per_step))))
        s2 = ClockSampler(local)
        if rank == 0:
            s2.start()
        barrier()
        rs = sess.run(0.0, n_sus)
        barrier()
        c2 = s2.stop() if rank == 0 else None
        ms_s = max_over_ranks(rs.device_ms)
        sustained = {"steps": n_sus, "device_seconds": ms_s * 1e-3, "ms_per_step": ms_s / n_sus, "value": R * 1e3 * n_sus / ms_s,
                     "unit": "iterations/s", "effective_tflops": flop_iter / (ms_s / n_sus * 1e-3) / 1e12, "clocks": c2}

    # ---- per-kernel timing for the roofline (eager run with cudaEvent pairs per phase) --------
    t = np.zeros(10)
    barrier()
    sess.run(0.0, 3, t=t)
    barrier()
    fused = sess.fused_steps()                                # bit 0: W step, bit 1: H step run as one kernel each
    corr = sess.corr_bf16()                                   # correction passes as BF16 MMAs: 8 instead of 12 MMA time units per block
    nq = 2 - bin(fused).count("1")                            # stand-alone quotient launches per iteration
    quot_ms = t[0] * 1e3 / (nq * 3) if nq else 0.0
    gemm2_ms = t[3] * 1e3 / (2 * 3)                           # the H-step and W-step GEMM kernels (fused: incl. their quotient)
    pk = peaks()
    if fused == 3:
        # dominant kernel = k_fused (quotient GEMM + consuming GEMM in one launch): 4*M*n_loc*K algorithmic flop per launch
        ach = 4.0 * M * n_loc * K * R / (gemm2_ms * 1e-3) / 1e12 if gemm2_ms > 0 else 0.0
        kernel = "k_fused<W step | H step> (W*H GEMM -> X./max(.,EPS) -> TF32 pair in smem -> second GEMM, 3xTF32)"
        launch_ms = gemm2_ms
        alg_bytes = 4.0 * M * n_loc + R * (8.0 * K * (M + n_loc) + 12.0 * K * min(M, n_loc))   # X read once; W, H pairs; the small output
    else:
        # dominant kernel = k_umma<QUOT> (W*H with the fused clamp/divide), 2*M*n_loc*K algorithmic flop per launch
        ach = 2.0 * M * n_loc * K / (quot_ms * 1e-3) / 1e12 if quot_ms > 0 else 0.0
        kernel = "k_umma<QUOT> (W*H GEMM + fused X./max(.,EPS), 3xTF32)"
        launch_ms = quot_ms
        alg_bytes = 8.0 * M * n_loc + 8.0 * K * (M + n_loc)   # X read + Y written once, W and H (hi, lo) read
    # ONE denominator, always: the MEASURED burst bf16 GEMM rate / 2 (TF32 runs at half the bf16 rate) / the TF32-pass
    # equivalents this kernel EXECUTES per fp32 product: 3 (hi*hi, lo*hi, hi*lo as TF32 MMAs), or 2 when the two correction
    # passes run as BF16 MMAs (each half a TF32 pass: 4 + 2 + 2 = 8 of 12 MMA time units per 32-deep block).
    # That measurement is itself a power-capped cuBLAS run, so a kernel that moves less through HBM can clock higher and
    # exceed it (frac > 1 is reported as such, never re-based).  The other views are separate keys; `frac_3xtf32` is against
    # the north star's 3xTF32 roofline (bf16 / 6) whatever the kernel executes.
    units = 1.0 if args.tf32_passes == 1 else (2.0 if corr else 3.0)
    peak = pk["bf16"] / 2.0 / units
    clk_mhz = (clocks or {}).get("sm_mhz") if rank == 0 else None
    max_mhz = float((clocks or {}).get("sm_max_mhz") or 1965.0)
    pipe = lambda mhz: 148 * 4096 * mhz * 1e6 / units / 1e12   # 148 SMs x 4096 dense TF32 flop/clk/SM, `units` TF32-pass equivalents per product
    traffic = None
    if world == 1:
        try:   # DRAM bytes per launch of the same kernel from the committed ncu --set full capture (profiles/traffic.json)
            tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
            traffic = tj.get(args.workload + ("+fused" if fused == 3 else ""), {}).get("1")
        except (OSError, ValueError):
            pass
    roofline = {"bound": "tensor", "kernel": kernel,
                "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak, "traffic": traffic,
                "traffic_source": "profiles/traffic.json (ncu --set full capture of this kernel and workload)" if traffic else None,
                "peak_note": f"{pk['source']}: burst bf16 {pk['bf16']:.1f} TF/s / 2 (TF32) / {units:g} (TF32-pass equivalents executed per product: "
                             + ("hi*hi as TF32 + two BF16 correction passes at half cost each" if units == 2.0 else "3xTF32") + "); fixed denominator",
                "executed_tf32_pass_equivalents": units, "corr_bf16": bool(corr),
                "frac_3xtf32": ach / (pk["bf16"] / 6.0),
                "frac_sustained": ach / (pk["bf16_sustained"] / 2.0 / units),
                "frac_nominal": ach / pipe(max_mhz),
                "frac_at_clock": ach / pipe(clk_mhz) if clk_mhz else None,
                "exceeds_measured_peak": bool(ach > peak),
                "algorithmic_hbm_bytes": alg_bytes,
                "launch_ms": launch_ms, "other_gemm_launch_ms": gemm2_ms if fused != 3 else None, "fused_steps": fused,
                "whole_iteration_frac": (flop_iter / world / (ms_per_step * 1e-3) / 1e12) / peak,   # same denominator
                "phase_seconds_3_iterations": [float(v) for v in t]}
    if world == 1:
        sess.close()      # N > 1: kept until the e2e run has shared its communicator (a process creates NCCL once)

    # ---- e2e: host buffers, H2D and D2H inside the timed region --------------------------------
    e2e = None
    if not args.no_e2e:
        g = torch.Generator(); g.manual_seed(seed * 1000 + rank)
        Xh = torch.empty((n_loc, M), dtype=torch.float32, pin_memory=True).uniform_(0, 1, generator=g)
        Whs = [torch.empty((K, M), dtype=torch.float32, pin_memory=True).uniform_(0, 1, generator=torch.Generator().manual_seed(seed + 7919 * r))
               for r in range(R)]
        Hhs = [torch.empty((n_loc, K), dtype=torch.float32, pin_memory=True).uniform_(0, 1, generator=g) for r in range(R)]
        Xn = Xh.numpy().T                                             # column-major views (M x n), (M x K), (K x n)
        Wns, Hns = [w.numpy().T for w in Whs], [h.numpy().T for h in Hhs]
        barrier()
        t0 = time.perf_counter()
        loop_ms = None
        if world == 1 and R == 1:
            r2 = api.update_div(Wns[0], Hns[0], Xn, 0.0, args.steps, opt=api.options(engine=args.engine, device=local, tf32_passes=args.tf32_passes))   # the reference-facing call
            loop_ms = r2.device_ms
            call = "update_div(W,H,X,0,steps)"
        elif world == 1:
            api.update_div_batch(Wns, Hns, Xn, 0.0, args.steps, opt=api.options(engine=args.engine, device=local))
            call = f"nmf_update_div_batch(W[{R}],H[{R}],X,0,steps)"
        else:
            s2 = api.Session(M, n_loc, K, n_offset=n_off, n_global=N, opt=opt)
            s2.share_comm(sess)                                            # the process-wide communicator
            s2.upload(Wns[0], Hns[0], Xn); r2 = s2.run(0.0, args.steps); s2.download(); s2.close()
            loop_ms = r2.device_ms
            call = "Session.upload/run/download per rank"
        torch.cuda.synchronize()
        dt = max_over_ranks(time.perf_counter() - t0)
        barrier()
        e2e = {"value": R * args.steps / dt, "unit": "iterations/s",
               "h2d_bytes_per_step": int(4 * (M * n_loc + R * (M * K + K * n_loc)) / args.steps),
               "d2h_bytes_per_step": int(4 * R * (M * K + K * n_loc) / args.steps),
               "seconds": dt, "call": call, "loop_device_ms": loop_ms,
               "outside_loop_seconds": (dt - loop_ms * 1e-3) if loop_ms else None,      # allocation + H2D + D2H (+ graph capture)
               "host_cpus_near_gpu": len(near_cpus) if near_cpus else None}
        # the box's own pinned-memory H2D rate, measured right here (1 GiB, best of 3): the call above is PCIe-bound, and the
        # boxes of the pool differ (52-54 GB/s on most, 12 GB/s seen on one) — this says which kind the line came from
        try:
            probe = Xh.view(-1)[: min(Xh.numel(), 1 << 28)]
            dst = torch.empty(probe.numel(), dtype=torch.float32, device=f"cuda:{local}")
            best = None
            for _ in range(3):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); dst.copy_(probe, non_blocking=True); e1.record(); e1.synchronize()
                ms = e0.elapsed_time(e1); best = ms if best is None else min(best, ms)
            e2e["h2d_probe_GBps"] = probe.numel() * 4 / (best * 1e-3) / 1e9
            e2e["h2d_call_GBps"] = (4.0 * M * n_loc / e2e["outside_loop_seconds"] / 1e9) if e2e.get("outside_loop_seconds") else None
            del dst
        except Exception as ex:                                   # a probe must never cost the line
            e2e["h2d_probe_GBps"] = None; e2e["h2d_probe_error"] = str(ex)[:100]
        del Xh, Whs, Hhs
    parity = None
    if world > 1:
        sess.close()
        parity = parity_vs_1gpu(api, torch, dist, rank, world, local)

    if rank == 0:
        xbytes = 4.0 * M * n_loc
        l2 = ("inputs (X alone is %.1f GB per GPU) far exceed the 126 MB L2; no flush needed" % (xbytes / 1e9) if xbytes > 2 * L2_BYTES
              else "warm L2: the whole working set (X %.1f MB per GPU) is L2-resident across iterations, as it is in a real run of this size" % (xbytes / 1e6))
        line = {"metric": "nmf_iterations_per_s", "value": value, "unit": "iterations/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
                "scaling": "weak" if weak else "strong", "vs_baseline": None, "dtype": ("tf32 (single-pass TF32 products, fp32 accumulate: reduced precision, NOT the headline arithmetic)" if args.tf32_passes == 1 else
                          "f32 (3xTF32 tensor-core products" + (", correction passes as BF16 MMAs" if corr else "") + ", fp32 accumulate)"),
                "data": "synthetic",
                "config": {"workload": desc, "M": M, "N": N, "K": K, "restarts": R, "columns_per_gpu": n_loc,
                           "parallelism": f"column-sharded x{world}", "engine": args.engine,
                           "hbm_resident_bytes_per_gpu": hbm_resident, "l2": l2},
                "effective_tflops": flop_iter / (ms_per_step * 1e-3) / 1e12,
                "gpu_launches": int(launches), "e2e": e2e, "roofline": roofline, "clocks": clocks, "sustained": sustained}
        if parity is not None:
            line["parity_vs_1gpu"] = parity
        if world == 1 and R == 1 and not args.no_refcuda:
            line["reference_cuda"] = reference_cuda(M, N, K, args.steps, args.warmup, local)
        if world == 1 and not args.no_cpu:
            if prev_affinity:
                os.sched_setaffinity(0, prev_affinity)         # the CPU baseline gets every host core
            line["cpu_baseline"] = cpu_baseline(M, N, K, seed)
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
