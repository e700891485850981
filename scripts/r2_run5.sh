#!/bin/bash
# round 2, GPU call 5: W-step tail rows: parity, restarts, then C2 / C3 with the tail path on / off
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x -p no:cacheprovider -k "tail_rows or restarts or batch or c2_ or c3_ or quotient_staging or multi or dbg or half_steps" > gpurun_out/r2_pytest5.log 2>&1; tail -8 gpurun_out/r2_pytest5.log
for tl in 1 0; do
for w in c2 c3; do
  NMF_W_TAIL=$tl timeout 600 python bench.py --workload $w --steps 50 --warmup 3 --no-cpu --no-e2e --no-refcuda --no-sustained > gpurun_out/r2_bench_${w}_tail$tl.json 2> gpurun_out/r2_bench_${w}_tail$tl.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r2_bench_${w}_tail$tl.json').read().strip().splitlines()[-1])
    print('$w tail=$tl', 'ms/iter', round(d['ms_per_step'],4), 'eff TF/s', round(d['effective_tflops'],1), 'launches/iter', d['gpu_launches']/50, 'phases us', [round(x/3*1e6,1) for x in d['roofline']['phase_seconds_3_iterations'][:6]], (d['clocks'] or {}).get('sm_mhz'))
except Exception as e:
    print('$w bench failed', e); print(open('gpurun_out/r2_bench_${w}_tail$tl.err').read()[-800:])
PY
done
done
