#!/bin/bash
# round 2, GPU call 18: woven stage order at K <= 128 (quotient in tensor memory) and on the small problems: A/B
mkdir -p gpurun_out
for rep in 1 2; do
for wv in 0 1; do
  for w in c3 c5k64 c2 c1 c2r8; do
  NMF_UMMA_WEAVE=$wv NMF_UMMA_DRAIN_EARLY=1 timeout 600 python bench.py --workload $w --steps 50 --warmup 5 --no-cpu --no-e2e --no-refcuda > gpurun_out/r2_bench_${w}_w$wv.json 2> gpurun_out/r2_bench_${w}_w$wv.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r2_bench_${w}_w$wv.json').read().strip().splitlines()[-1]); r=d['roofline']
    print('$w weave=$wv', 'ms/iter', round(d['ms_per_step'],5), 'sustained', round(d['sustained']['ms_per_step'],5), 'eff TF/s', round(d['effective_tflops'],1), 'frac', round(r['frac'],3), (d['clocks'] or {}).get('sm_mhz'))
except Exception as e:
    print('$w weave=$wv bench failed', e); print(open('gpurun_out/r2_bench_${w}_w$wv.err').read()[-800:])
PY
  done
done
done
