#!/bin/bash
# round 2, GPU call 21: single-pass TF32 accuracy test + bench line; the driver's round-end commands
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -s -p no:cacheprovider -k "single_pass" 2>&1 | grep -E "single-pass|passed|failed|Error" | head
for w in c4 c3; do
timeout 600 python bench.py --workload $w --steps 20 --warmup 3 --tf32-passes 1 --no-cpu --no-refcuda > gpurun_out/r2_bench_${w}_tf32x1.json 2> gpurun_out/r2_bench_${w}_tf32x1.err
python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r2_bench_${w}_tf32x1.json').read().strip().splitlines()[-1]); r=d['roofline']
    print('$w tf32x1', 'ms/iter', round(d['ms_per_step'],4), 'sustained', round(d['sustained']['ms_per_step'],4), 'eff TF/s', round(d['effective_tflops'],1), 'frac', round(r['frac'],3), 'fused', r['fused_steps'], d['dtype'][:40], (d['clocks'] or {}).get('sm_mhz'))
except Exception as e:
    print('$w tf32x1 bench failed', e); print(open('gpurun_out/r2_bench_${w}_tf32x1.err').read()[-800:])
PY
done
timeout 1800 python -m pytest tests/ -x -q -m gpu -p no:cacheprovider > gpurun_out/r2_pytest_gpu.log 2>&1; tail -3 gpurun_out/r2_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 600 python bench.py --impl reference --gpus 1 --steps 2 --warmup 1 2>/dev/null | cut -c1-300
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2_bench_c4_n1.json 2> gpurun_out/r2_bench_c4_n1.err; python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2_bench_c4_n1.json').read().strip().splitlines()[-1]); r=d['roofline']
print('c4 default', d['ms_per_step'], d['value'], d['e2e']['value'], r['frac'], r['frac_3xtf32'], d['clocks'], d['gpu_launches'])
PY
