#!/bin/bash
# round 2, GPU call 2: every BASELINE workload on one GPU (with the reference's CUDA build timed next to it), the restarts
# workloads, and a full ncu capture of the K = 64 fused kernels (the ones furthest below the tensor roofline)
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x -p no:cacheprovider > gpurun_out/r2_pytest2.log 2>&1; tail -15 gpurun_out/r2_pytest2.log
for w in c1 c2 c2r4 c2r8 c3 c5k64 c5 c5k512; do
  timeout 900 python bench.py --workload $w --steps 50 --warmup 3 --no-cpu > gpurun_out/r2_bench_$w.json 2> gpurun_out/r2_bench_$w.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r2_bench_$w.json').read().strip().splitlines()[-1])
    rc=d.get('reference_cuda') or {}
    print('$w', 'ms/iter', round(d['ms_per_step'],4), 'value', round(d['value'],2), 'eff TF/s', round(d['effective_tflops'],1), 'fused', d['roofline']['fused_steps'],
          'launches/iter', d['gpu_launches']/50, 'sustained ms', round(d['sustained']['ms_per_step'],4), 'e2e', round(d['e2e']['value'],2),
          'ref_cuda ms', rc.get('ms_per_step', rc.get('unavailable')), 'phases us', [round(x/3*1e6,1) for x in d['roofline']['phase_seconds_3_iterations'][:6]], (d['clocks'] or {}).get('sm_mhz'))
except Exception as e:
    print('$w bench failed', e); print(open('gpurun_out/r2_bench_$w.err').read()[-800:])
PY
done
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_fused -s 6 -c 2 -o gpurun_out/r2_fused_c5k64 -f python bench.py --workload c5k64 --steps 2 --warmup 3 --no-e2e --no-cpu --no-refcuda --no-sustained > gpurun_out/ncu_full_c5k64.log 2>&1; echo "ncu c5k64 rc=$?"
ls -la gpurun_out/*.ncu-rep
