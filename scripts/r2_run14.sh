#!/bin/bash
# round 2, GPU call 14: the BF16-correction kernels are bound by the shared-memory data pipe (tensor-core operand reads + TMA
# writes + worker stores ~ 90 % of 128 B/clk): one 256-wide S tile halves the a-stage A reads and the re-streamed rows — A/B
mkdir -p gpurun_out
for cfg in "128 0" "256 0" "256 1" "128 0" "256 0"; do
  set -- $cfg
  for w in c5 c4; do
  NMF_UMMA_FUSE_BN=$1 NMF_UMMA_FUSE_ALT=$2 timeout 600 python bench.py --workload $w --steps 20 --warmup 3 --no-cpu --no-e2e --no-refcuda > gpurun_out/r2_bench_${w}_bn$1_alt$2.json 2> gpurun_out/r2_bench_${w}_bn$1_alt$2.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r2_bench_${w}_bn$1_alt$2.json').read().strip().splitlines()[-1]); r=d['roofline']
    print('$w bn=$1 alt=$2', 'ms/iter', round(d['ms_per_step'],4), 'sustained', round(d['sustained']['ms_per_step'],4), 'eff TF/s', round(d['effective_tflops'],1), 'frac', round(r['frac'],3), 'phases us', [round(x/3*1e6,1) for x in r['phase_seconds_3_iterations'][:8]], (d['clocks'] or {}).get('sm_mhz'))
except Exception as e:
    print('$w bn=$1 alt=$2 bench failed', e); print(open('gpurun_out/r2_bench_${w}_bn$1_alt$2.err').read()[-800:])
PY
  done
done
NMF_UMMA_FUSE_BN=256 timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_fused -s 6 -c 2 -o gpurun_out/r2_fused_c4_bn256 -f python bench.py --workload c4 --steps 2 --warmup 3 --no-e2e --no-cpu --no-refcuda --no-sustained > gpurun_out/ncu_full_c4_bn256.log 2>&1; echo "ncu c4 bn256 rc=$?"
