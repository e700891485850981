#!/bin/bash
# round 2, GPU call 17: woven a/b stage order (NMF_UMMA_WEAVE): parity, then A/B at C4, C5 with the other switches
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -p no:cacheprovider -k "alternating" > gpurun_out/r2_pytest17.log 2>&1; tail -5 gpurun_out/r2_pytest17.log
for cfg in "0 0 0" "0 0 1" "0 1 1" "1 1 1" "1 1 0" "0 1 1" "1 1 1"; do
  set -- $cfg
  for w in c5 c4; do
  NMF_UMMA_FUSE_ALT=$1 NMF_UMMA_DRAIN_EARLY=$2 NMF_UMMA_WEAVE=$3 timeout 600 python bench.py --workload $w --steps 20 --warmup 3 --no-cpu --no-e2e --no-refcuda > gpurun_out/r2_bench_${w}_a$1e$2w$3.json 2> gpurun_out/r2_bench_${w}_a$1e$2w$3.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r2_bench_${w}_a$1e$2w$3.json').read().strip().splitlines()[-1]); r=d['roofline']
    print('$w alt=$1 early=$2 weave=$3', 'ms/iter', round(d['ms_per_step'],4), 'sustained', round(d['sustained']['ms_per_step'],4), 'eff TF/s', round(d['effective_tflops'],1), 'frac', round(r['frac'],3), (d['clocks'] or {}).get('sm_mhz'))
except Exception as e:
    print('$w alt=$1 early=$2 weave=$3 bench failed', e); print(open('gpurun_out/r2_bench_${w}_a$1e$2w$3.err').read()[-800:])
PY
  done
done
NMF_UMMA_FUSE_ALT=1 NMF_UMMA_DRAIN_EARLY=1 NMF_UMMA_WEAVE=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_fused -s 6 -c 2 -o gpurun_out/r2_fused_c4_weave -f python bench.py --workload c4 --steps 2 --warmup 3 --no-e2e --no-cpu --no-refcuda --no-sustained > gpurun_out/ncu_full_c4_weave.log 2>&1; echo "ncu c4 weave rc=$?"
