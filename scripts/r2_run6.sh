#!/bin/bash
# round 2, GPU call 6: whole GPU suite, every workload's bench line (reference CUDA build next to ours), HBM evidence for the helper kernels
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider --durations=8 > gpurun_out/r2_pytest6.log 2>&1; tail -14 gpurun_out/r2_pytest6.log
for w in c1 c2 c2r4 c2r8 c3 c5k64 c5 c5k512; do
  timeout 900 python bench.py --workload $w --steps 50 --warmup 3 --no-cpu > gpurun_out/r2_bench_$w.json 2> gpurun_out/r2_bench_$w.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r2_bench_$w.json').read().strip().splitlines()[-1])
    rc=d.get('reference_cuda') or {}
    print('$w', 'ms/iter', round(d['ms_per_step'],4), 'value', round(d['value'],2), 'eff TF/s', round(d['effective_tflops'],1), 'fused', d['roofline']['fused_steps'],
          'launches/iter', d['gpu_launches']/50, 'sustained ms', round(d['sustained']['ms_per_step'],4), 'e2e', round(d['e2e']['value'],2),
          'ref_cuda ms', rc.get('ms_per_step', rc.get('unavailable')), 'phases us', [round(x/3*1e6,1) for x in d['roofline']['phase_seconds_3_iterations'][:6]], (d['clocks'] or {}).get('sm_mhz'))
except Exception as e:
    print('$w bench failed', e); print(open('gpurun_out/r2_bench_$w.err').read()[-800:])
PY
done
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench_c4_n1.json 2> gpurun_out/r2_bench_c4_n1.err; echo "c4 rc=$?"; cut -c1-400 gpurun_out/r2_bench_c4_n1.json
timeout 300 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r2_bench_reference_arm.json 2>&1; cut -c1-300 gpurun_out/r2_bench_reference_arm.json
for w in c4 c3; do
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,dram__bytes_read.sum.per_second,dram__bytes_write.sum.per_second --clock-control none -k regex:'k_colsum|k_rowsum|k_update_w|k_update_h|k_w_tail|k_fold|k_sum' -c 40 --csv --log-file gpurun_out/r2_hbm_helpers_$w.csv python bench.py --workload $w --steps 2 --warmup 3 --no-e2e --no-cpu --no-refcuda --no-sustained > gpurun_out/ncu_helpers_$w.log 2>&1; echo "ncu helpers $w rc=$?"
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches_c4.csv python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu --no-refcuda --no-sustained > gpurun_out/ncu_launch_c4.log 2>&1; echo "ncu launches rc=$?"
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
