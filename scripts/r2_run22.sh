#!/bin/bash
# round 2, GPU call 22: BF16 correction passes in the two-kernel path (K > 256, NMF_CORR_BF16 bit 2): parity, A/B at C5 K=512
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -p no:cacheprovider -k "two_kernel_bf16" > gpurun_out/r2_pytest22.log 2>&1; tail -25 gpurun_out/r2_pytest22.log | cut -c1-300
for rep in 1 2; do
for c in 3 7; do
  NMF_CORR_BF16=$c timeout 600 python bench.py --workload c5k512 --steps 30 --warmup 3 --no-cpu --no-e2e --no-refcuda > gpurun_out/r2_bench_c5k512_corr$c.json 2> gpurun_out/r2_bench_c5k512_corr$c.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r2_bench_c5k512_corr$c.json').read().strip().splitlines()[-1]); r=d['roofline']
    print('c5k512 corr=$c', 'ms/iter', round(d['ms_per_step'],4), 'sustained', round(d['sustained']['ms_per_step'],4), 'eff TF/s', round(d['effective_tflops'],1), 'frac', round(r['frac'],3), 'corr', r['corr_bf16'], 'phases us', [round(x/3*1e6,1) for x in r['phase_seconds_3_iterations'][:8]], (d['clocks'] or {}).get('sm_mhz'))
except Exception as e:
    print('c5k512 corr=$c bench failed', e); print(open('gpurun_out/r2_bench_c5k512_corr$c.err').read()[-800:])
PY
done
done
