#!/bin/bash
# round 2, GPU call 3: the quotient staged in tensor memory (K <= 128): parity, then K <= 128 workloads with it on / off
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -p no:cacheprovider -k "quotient_staging or fused_kernel_variants or c1_exact or fused_iterations_c1 or extreme_scales or fused_h_step or fused_w_step" > gpurun_out/r2_pytest3.log 2>&1; tail -15 gpurun_out/r2_pytest3.log
for yt in 1 0; do
for w in c1 c2 c3 c5k64; do
  NMF_UMMA_FUSE_YT=$yt timeout 600 python bench.py --workload $w --steps 50 --warmup 3 --no-cpu --no-e2e --no-refcuda --no-sustained > gpurun_out/r2_bench_${w}_yt$yt.json 2> gpurun_out/r2_bench_${w}_yt$yt.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r2_bench_${w}_yt$yt.json').read().strip().splitlines()[-1])
    print('$w yt=$yt', 'ms/iter', round(d['ms_per_step'],4), 'eff TF/s', round(d['effective_tflops'],1), 'phases us', [round(x/3*1e6,1) for x in d['roofline']['phase_seconds_3_iterations'][:6]], (d['clocks'] or {}).get('sm_mhz'))
except Exception as e:
    print('$w bench failed', e); print(open('gpurun_out/r2_bench_${w}_yt$yt.err').read()[-800:])
PY
done
done
