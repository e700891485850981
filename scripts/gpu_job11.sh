timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_gpu.log | cut -c1-300
bash scripts/ref_cuda_c1.sh 2>&1 | grep -v "^+" 
for w in c1 c2; do
timeout 600 python bench.py --workload $w --steps 50 --warmup 3 --no-cpu --no-e2e > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; python - <<PY
import json
d=json.loads(open('gpurun_out/bench_$w.json').read().strip().splitlines()[-1])
print('$w', 'ms/iter', round(d['ms_per_step'],4), 'phases', [round(x/3*1e6,1) for x in d['roofline']['phase_seconds_3_iterations']])
PY
done
