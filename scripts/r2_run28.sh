#!/bin/bash
# round 2, GPU call 28: the bench line with the H2D probe (default C4 line + C5)
mkdir -p gpurun_out
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2_bench_c4_n1.json 2> gpurun_out/r2_bench_c4_n1.err
timeout 600 python bench.py --workload c5 --steps 50 --warmup 5 > gpurun_out/r2_bench_c5.json 2> gpurun_out/r2_bench_c5.err
python - <<'PY'
import json
for w in ('c4_n1','c5'):
    try:
        d=json.loads(open(f'gpurun_out/r2_bench_{w}.json').read().strip().splitlines()[-1]); r=d['roofline']
        print(w, d['ms_per_step'], d['value'], 'e2e', d['e2e'], 'frac', r['frac'], r['frac_3xtf32'], d['clocks'], (d.get('reference_cuda') or {}).get('ms_per_step'))
    except Exception as e:
        print(w, 'failed', e); print(open(f'gpurun_out/r2_bench_{w}.err').read()[-1500:])
PY
