set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_gpu.log | cut -c1-300
export NMF_UMMA_CHUNK=4
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e > gpurun_out/bench_a.json 2>&1; python - <<PY
import json
d=json.loads(open('gpurun_out/bench_a.json').read().strip().splitlines()[-1])
print('bench', 'ms/iter', round(d['ms_per_step'],2), 'G1', round(d['roofline']['launch_ms'],2), 'G2/3 avg', round(d['roofline']['other_gemm_launch_ms'],2), d['clocks'])
PY
timeout 600 python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/plain.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/launches_b.csv python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_umma -s 8 -c 4 -o gpurun_out/umma_b_c4 -f python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_full_c4.log 2>&1; echo "ncu full c4 rc=$?"
