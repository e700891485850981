#!/bin/bash
# round 2, GPU call 12: L2 -> SM delivery microbenchmark (the bound of the BF16-correction kernels); full GPU suite with the
# BF16 corrections on at K = 128 too (new default); C1 / C3 lines; ncu capture of the C3 kernels
mkdir -p gpurun_out
B=nmf_gpu_b200/build/l2_to_sm_bw
for a in "48 16 4 1" "48 16 8 1" "48 32 4 1" "48 16 4 2" "48 8 8 2" "16 16 8 1" "96 16 8 1" "4096 16 8 1"; do timeout 120 $B $a; done | tee gpurun_out/r2_l2_to_sm_bw.jsonl
timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/r2_pytest_gpu.log 2>&1; tail -5 gpurun_out/r2_pytest_gpu.log
for w in c1 c3; do
  timeout 900 python bench.py --workload $w --steps 50 --warmup 5 > gpurun_out/r2_bench_$w.json 2> gpurun_out/r2_bench_$w.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r2_bench_$w.json').read().strip().splitlines()[-1]); r=d['roofline']
    print('$w', 'ms/iter', round(d['ms_per_step'],4), 'sustained', round(d['sustained']['ms_per_step'],4), 'eff TF/s', round(d['effective_tflops'],1), 'frac', round(r['frac'],3), 'frac_3xtf32', round(r['frac_3xtf32'],3), 'e2e', d['e2e']['value'], 'refcuda', (d.get('reference_cuda') or {}).get('ms_per_iteration'), 'corr', r['corr_bf16'])
except Exception as e:
    print('$w bench failed', e); print(open('gpurun_out/r2_bench_$w.err').read()[-800:])
PY
done
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_fused -s 6 -c 2 -o gpurun_out/r2_fused_c3_cb -f python bench.py --workload c3 --steps 2 --warmup 3 --no-e2e --no-cpu --no-refcuda --no-sustained > gpurun_out/ncu_full_c3_cb.log 2>&1; echo "ncu c3 rc=$?"
