timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_gpu.log | cut -c1-300
for w in c4 c3 c5k64 c2 c1; do
timeout 600 python bench.py --workload $w --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/bench_$w.json').read().strip().splitlines()[-1])
    print('$w', 'ms/iter', round(d['ms_per_step'],4), 'eff TF/s', round(d['effective_tflops'],1), 'G1', round(d['roofline']['launch_ms'],3), 'G2/3 avg', round(d['roofline']['other_gemm_launch_ms'],3), d['clocks']['sm_mhz'])
except Exception as e:
    print('$w bench failed', e); print(open('gpurun_out/bench_$w.err').read()[-800:])
PY
done
