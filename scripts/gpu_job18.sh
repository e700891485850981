for rep in 1 2; do
for hh in 0 2 1; do
NMF_UMMA_HINT_H=$hh timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e > gpurun_out/bench_ab.json 2>&1; python - <<PY
import json
d=json.loads(open('gpurun_out/bench_ab.json').read().strip().splitlines()[-1])
print('hint_h $hh', 'ms/iter', round(d['ms_per_step'],2), 'G1', round(d['roofline']['launch_ms'],2), 'G2/3 avg', round(d['roofline']['other_gemm_launch_ms'],2), d['clocks']['sm_mhz'])
PY
done
done
