for w in c1 c2 c3 c5k64 c5 c5k512; do
timeout 600 python bench.py --workload $w --steps 20 --warmup 3 --no-cpu > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/bench_$w.json').read().strip().splitlines()[-1])
    print('$w', 'ms/iter', round(d['ms_per_step'],3), 'eff TF/s', round(d['effective_tflops'],1), 'G1', round(d['roofline']['launch_ms'],3), 'G2/3 avg', round(d['roofline']['other_gemm_launch_ms'],3), 'e2e it/s', round(d['e2e']['value'],2), 'phases', [round(x,5) for x in d['roofline']['phase_seconds_3_iterations']])
except Exception as e:
    print('$w bench failed', e); print(open('gpurun_out/bench_$w.err').read()[-800:])
PY
done
