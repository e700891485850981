#!/bin/bash
# round 2, GPU call 24: pre-split quotient storage in the K > 256 two-kernel path: parity, A/B at C5 K=512
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_headline.py -m gpu -q -x -p no:cacheprovider -k "two_kernel_bf16 or k512 or quotient" > gpurun_out/r2_pytest24.log 2>&1; tail -25 gpurun_out/r2_pytest24.log | cut -c1-300
for rep in 1 2; do
for c in 0 1; do
  NMF_UMMA_PRESPLIT=$c timeout 600 python bench.py --workload c5k512 --steps 30 --warmup 3 --no-cpu --no-e2e --no-refcuda > gpurun_out/r2_bench_c5k512_pre$c.json 2> gpurun_out/r2_bench_c5k512_pre$c.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r2_bench_c5k512_pre$c.json').read().strip().splitlines()[-1]); r=d['roofline']
    print('c5k512 presplit=$c', 'ms/iter', round(d['ms_per_step'],4), 'sustained', round(d['sustained']['ms_per_step'],4), 'eff TF/s', round(d['effective_tflops'],1), 'frac', round(r['frac'],3), 'phases us', [round(x/3*1e6,1) for x in r['phase_seconds_3_iterations'][:8]], (d['clocks'] or {}).get('sm_mhz'))
except Exception as e:
    print('c5k512 presplit=$c bench failed', e); print(open('gpurun_out/r2_bench_c5k512_pre$c.err').read()[-800:])
PY
done
done
