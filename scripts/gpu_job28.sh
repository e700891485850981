NMF_B200_LIB=$PWD/nmf_gpu_b200/libnmf_b200_trunc.so timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "raw_gemm or half_steps or c1_exact or chunk" 2>&1 | tail -5 | cut -c1-200
NMF_B200_LIB=$PWD/nmf_gpu_b200/libnmf_b200_trunc.so python scripts/bias_probe.py
for rep in 1 2; do
for lib in libnmf_b200.so libnmf_b200_trunc.so; do
NMF_B200_LIB=$PWD/nmf_gpu_b200/$lib timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e > gpurun_out/bench_ab.json 2>&1; python - <<PY
import json
d=json.loads(open('gpurun_out/bench_ab.json').read().strip().splitlines()[-1])
print('$lib', 'ms/iter', round(d['ms_per_step'],2), 'G1', round(d['roofline']['launch_ms'],2), 'G2/3 avg', round(d['roofline']['other_gemm_launch_ms'],2), d['clocks']['sm_mhz'])
PY
done
done
