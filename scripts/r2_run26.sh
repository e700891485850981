#!/bin/bash
# round 2, GPU call 26: 12 epilogue warps in the quotient kernel at K = 512 now that its MMAs take a third less (NMF_UMMA_W3_MAXK)
mkdir -p gpurun_out
for rep in 1 2; do
for k in 192 512; do
  NMF_UMMA_W3_MAXK=$k timeout 600 python bench.py --workload c5k512 --steps 30 --warmup 3 --no-cpu --no-e2e --no-refcuda > gpurun_out/r2_bench_c5k512_w3k$k.json 2> gpurun_out/r2_bench_c5k512_w3k$k.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r2_bench_c5k512_w3k$k.json').read().strip().splitlines()[-1]); r=d['roofline']
    print('c5k512 w3_max_k=$k', 'ms/iter', round(d['ms_per_step'],4), 'sustained', round(d['sustained']['ms_per_step'],4), 'eff TF/s', round(d['effective_tflops'],1), 'phases us', [round(x/3*1e6,1) for x in r['phase_seconds_3_iterations'][:8]], (d['clocks'] or {}).get('sm_mhz'))
except Exception as e:
    print('c5k512 w3_max_k=$k bench failed', e); print(open('gpurun_out/r2_bench_c5k512_w3k$k.err').read()[-800:])
PY
done
done
