#!/bin/bash
# round 2, GPU call 23: BF16 correction passes on by default in the two-kernel path too (K > 256): whole GPU suite, C5 K=512 lines
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/r2_pytest_gpu.log 2>&1; tail -4 gpurun_out/r2_pytest_gpu.log
for c in 7 3 7; do
  NMF_CORR_BF16=$c timeout 600 python bench.py --workload c5k512 --steps 30 --warmup 3 --no-cpu > gpurun_out/r2_bench_c5k512_corr$c.json 2> gpurun_out/r2_bench_c5k512_corr$c.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r2_bench_c5k512_corr$c.json').read().strip().splitlines()[-1]); r=d['roofline']
    print('c5k512 corr=$c', 'ms/iter', round(d['ms_per_step'],4), 'sustained', round(d['sustained']['ms_per_step'],4), 'eff TF/s', round(d['effective_tflops'],1), 'frac', round(r['frac'],3), 'frac3x', round(r['frac_3xtf32'],3), 'corr', r['corr_bf16'], 'e2e', round(d['e2e']['value'],1), 'ref', (d.get('reference_cuda') or {}).get('ms_per_step'), 'phases us', [round(x/3*1e6,1) for x in r['phase_seconds_3_iterations'][:8]], (d['clocks'] or {}).get('sm_mhz'))
except Exception as e:
    print('c5k512 corr=$c bench failed', e); print(open('gpurun_out/r2_bench_c5k512_corr$c.err').read()[-800:])
PY
done
