#!/bin/bash
# round 2, GPU call 8: after the L2 prefetch of H in the fused H step: parity subset, K <= 128 workloads, C3 as specified
# (convergence run), and a full capture of the fused kernels on one rank's shard of the 8-GPU run (the N = 8 limiter)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x -p no:cacheprovider -k "fused or c1_exact or c3_ or tail or restarts or batch" > gpurun_out/r2_pytest8.log 2>&1; tail -4 gpurun_out/r2_pytest8.log
for w in c3 c5k64 c5 c4s8; do
  timeout 600 python bench.py --workload $w --steps 50 --warmup 3 --no-cpu --no-e2e --no-refcuda > gpurun_out/r2_bench_${w}_pf.json 2> gpurun_out/r2_bench_${w}_pf.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r2_bench_${w}_pf.json').read().strip().splitlines()[-1])
    print('$w', 'ms/iter', round(d['ms_per_step'],4), 'sustained', round(d['sustained']['ms_per_step'],4), 'eff TF/s', round(d['effective_tflops'],1), 'launches/iter', d['gpu_launches']/50, 'phases us', [round(x/3*1e6,1) for x in d['roofline']['phase_seconds_3_iterations'][:8]], (d['clocks'] or {}).get('sm_mhz'))
except Exception as e:
    print('$w bench failed', e); print(open('gpurun_out/r2_bench_${w}_pf.err').read()[-800:])
PY
done
timeout 900 python scripts/run_c3_convergence.py > gpurun_out/r2_c3_convergence.json 2> gpurun_out/r2_c3_convergence.err; cut -c1-600 gpurun_out/r2_c3_convergence.json; tail -2 gpurun_out/r2_c3_convergence.err
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_fused -s 6 -c 2 -o gpurun_out/r2_fused_c4s8 -f python bench.py --workload c4s8 --steps 2 --warmup 3 --no-e2e --no-cpu --no-refcuda --no-sustained > gpurun_out/ncu_full_c4s8.log 2>&1; echo "ncu c4s8 rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r2_launches_c4s8.csv python bench.py --workload c4s8 --steps 2 --warmup 3 --no-e2e --no-cpu --no-refcuda --no-sustained > /dev/null 2>&1; echo "ncu launches c4s8 rc=$?"
