#!/bin/bash
# round 2, GPU call 19: the whole GPU suite and every single-GPU bench line with the round's final defaults
# (BF16 correction passes at K = 256 and 128, woven stage order + early drain at K > 128)
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/r2_pytest_gpu.log 2>&1; tail -5 gpurun_out/r2_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
for w in c4 c5 c3 c5k64 c5k512 c2 c1 c2r4 c2r8; do
  steps=50; [ $w = c4 ] && steps=20
  timeout 900 python bench.py --workload $w --steps $steps --warmup 5 > gpurun_out/r2_bench_$w.json 2> gpurun_out/r2_bench_$w.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r2_bench_$w.json').read().strip().splitlines()[-1]); r=d['roofline']
    print('$w', 'ms/iter', round(d['ms_per_step'],5), 'sustained', round(d['sustained']['ms_per_step'],5), 'it/s', round(d['value'],2), 'eff TF/s', round(d['effective_tflops'],1), 'frac', round(r['frac'],3), 'frac_3xtf32', round(r['frac_3xtf32'],3), 'e2e', round(d['e2e']['value'],2), 'refcuda ms', (d.get('reference_cuda') or {}).get('ms_per_step'), 'cpu', (d.get('cpu_baseline') or {}).get('value'), (d['clocks'] or {}).get('sm_mhz'), (d['sustained']['clocks'] or {}).get('sm_mhz'))
except Exception as e:
    print('$w bench failed', e); print(open('gpurun_out/r2_bench_$w.err').read()[-800:])
PY
done
cp gpurun_out/r2_bench_c4.json gpurun_out/r2_bench_c4_n1.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2_bench_reference_arm.json 2> gpurun_out/r2_bench_reference_arm.err; cut -c1-400 gpurun_out/r2_bench_reference_arm.json
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_fused -s 6 -c 2 -o gpurun_out/r2_fused_c4_final -f python bench.py --workload c4 --steps 2 --warmup 3 --no-e2e --no-cpu --no-refcuda --no-sustained > gpurun_out/ncu_full_c4_final.log 2>&1; echo "ncu c4 rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r2_launches_c4.csv python bench.py --workload c4 --steps 2 --warmup 3 --no-e2e --no-cpu --no-refcuda --no-sustained > /dev/null 2>&1; echo "ncu launches c4 rc=$?"
