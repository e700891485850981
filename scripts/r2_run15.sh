#!/bin/bash
# round 2, GPU call 15: early drain of P (before the ring-slot wait of a tile's first chunk) — parity, then A/B at C4, C5
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -p no:cacheprovider -k "alternating" > gpurun_out/r2_pytest15.log 2>&1; tail -5 gpurun_out/r2_pytest15.log
for cfg in "0 128" "1 128" "1 256" "0 128" "1 128"; do
  set -- $cfg
  for w in c5 c4; do
  NMF_UMMA_DRAIN_EARLY=$1 NMF_UMMA_FUSE_BN=$2 timeout 600 python bench.py --workload $w --steps 20 --warmup 3 --no-cpu --no-e2e --no-refcuda > gpurun_out/r2_bench_${w}_early$1_bn$2.json 2> gpurun_out/r2_bench_${w}_early$1_bn$2.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r2_bench_${w}_early$1_bn$2.json').read().strip().splitlines()[-1]); r=d['roofline']
    print('$w early=$1 bn=$2', 'ms/iter', round(d['ms_per_step'],4), 'sustained', round(d['sustained']['ms_per_step'],4), 'eff TF/s', round(d['effective_tflops'],1), 'frac', round(r['frac'],3), (d['clocks'] or {}).get('sm_mhz'))
except Exception as e:
    print('$w early=$1 bn=$2 bench failed', e); print(open('gpurun_out/r2_bench_${w}_early$1_bn$2.err').read()[-800:])
PY
  done
done
NMF_UMMA_DRAIN_EARLY=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_fused -s 6 -c 2 -o gpurun_out/r2_fused_c4_early -f python bench.py --workload c4 --steps 2 --warmup 3 --no-e2e --no-cpu --no-refcuda --no-sustained > gpurun_out/ncu_full_c4_early.log 2>&1; echo "ncu c4 early rc=$?"
