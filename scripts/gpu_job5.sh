set -x
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "raw_gemm" > gpurun_out/pytest_raw.log 2>&1; echo "raw rc=$?"; tail -5 gpurun_out/pytest_raw.log | cut -c1-300
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_gpu.log | cut -c1-300
for c in 2 4; do
NMF_UMMA_CHUNK=$c timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e > gpurun_out/bench_chunk$c.json 2>&1; python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/bench_chunk$c.json').read().strip().splitlines()[-1])
    print('chunk $c', 'ms/iter', round(d['ms_per_step'],2), 'G1', round(d['roofline']['launch_ms'],2), 'G2/3 avg', round(d['roofline']['other_gemm_launch_ms'],2), d['clocks'])
except Exception as e:
    print('bench failed', e); print(open('gpurun_out/bench_chunk$c.json').read()[-1500:])
PY
done
