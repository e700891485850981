timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log | cut -c1-300
for cap in 256 128; do
for w in c3 c5k64 c2; do
NMF_UMMA_NP_BNCAP=$cap timeout 600 python bench.py --workload $w --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; python - <<PY
import json
d=json.loads(open('gpurun_out/bench_$w.json').read().strip().splitlines()[-1])
print('cap $cap', '$w', 'ms/iter', round(d['ms_per_step'],4), 'G1', round(d['roofline']['launch_ms'],3), 'G2/3 avg', round(d['roofline']['other_gemm_launch_ms'],3), d['clocks']['sm_mhz'])
PY
done
done
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e > gpurun_out/bench_c4.json 2>&1;  python - <<PY
import json
d=json.loads(open('gpurun_out/bench_c4.json').read().strip().splitlines()[-1])
print('c4', 'ms/iter', round(d['ms_per_step'],4), 'G1', round(d['roofline']['launch_ms'],3), 'G2/3 avg', round(d['roofline']['other_gemm_launch_ms'],3), d['clocks']['sm_mhz'])
PY
