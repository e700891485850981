#!/bin/bash
# round 2, GPU call 9: BF16 correction passes (K = 256): parity in its own process (a wrong descriptor can trap), then A/B
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -p no:cacheprovider -k "bf16_correction" > gpurun_out/r2_pytest9.log 2>&1; tail -30 gpurun_out/r2_pytest9.log | cut -c1-300
for cb in 1 0; do
for w in c5 c4; do
  NMF_CORR_BF16=$cb timeout 600 python bench.py --workload $w --steps 20 --warmup 3 --no-cpu --no-e2e --no-refcuda > gpurun_out/r2_bench_${w}_cb$cb.json 2> gpurun_out/r2_bench_${w}_cb$cb.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r2_bench_${w}_cb$cb.json').read().strip().splitlines()[-1])
    print('$w corr_bf16=$cb', 'ms/iter', round(d['ms_per_step'],4), 'sustained', round(d['sustained']['ms_per_step'],4), (d['sustained']['clocks'] or {}).get('sm_mhz'), 'eff TF/s', round(d['effective_tflops'],1), 'phases us', [round(x/3*1e6,1) for x in d['roofline']['phase_seconds_3_iterations'][:8]], (d['clocks'] or {}).get('sm_mhz'))
except Exception as e:
    print('$w bench failed', e); print(open('gpurun_out/r2_bench_${w}_cb$cb.err').read()[-800:])
PY
done
done
