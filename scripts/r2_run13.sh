#!/bin/bash
# round 2, GPU call 13: alternating chunk assignment of the K > 128 worker warps (NMF_UMMA_FUSE_ALT): parity, then A/B at C4, C5
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -p no:cacheprovider -k "alternating" > gpurun_out/r2_pytest13.log 2>&1; tail -5 gpurun_out/r2_pytest13.log
for rep in 1 2; do
for a in 0 1; do
  for w in c5 c4; do
  NMF_UMMA_FUSE_ALT=$a timeout 600 python bench.py --workload $w --steps 20 --warmup 3 --no-cpu --no-e2e --no-refcuda > gpurun_out/r2_bench_${w}_alt$a.json 2> gpurun_out/r2_bench_${w}_alt$a.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r2_bench_${w}_alt$a.json').read().strip().splitlines()[-1]); r=d['roofline']
    print('$w alt=$a', 'ms/iter', round(d['ms_per_step'],4), 'sustained', round(d['sustained']['ms_per_step'],4), 'eff TF/s', round(d['effective_tflops'],1), 'frac', round(r['frac'],3), 'phases us', [round(x/3*1e6,1) for x in r['phase_seconds_3_iterations'][:8]], (d['clocks'] or {}).get('sm_mhz'))
except Exception as e:
    print('$w alt=$a bench failed', e); print(open('gpurun_out/r2_bench_${w}_alt$a.err').read()[-800:])
PY
  done
done
done
NMF_UMMA_FUSE_ALT=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_fused -s 6 -c 2 -o gpurun_out/r2_fused_c4_alt -f python bench.py --workload c4 --steps 2 --warmup 3 --no-e2e --no-cpu --no-refcuda --no-sustained > gpurun_out/ncu_full_c4_alt.log 2>&1; echo "ncu c4 alt rc=$?"
