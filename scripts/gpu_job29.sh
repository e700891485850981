for rep in 1 2; do
for x in 0 1; do
for w in c3 c5k64 c4; do
NMF_UMMA_XPF=$x timeout 600 python bench.py --workload $w --steps 10 --warmup 3 --no-cpu --no-e2e > gpurun_out/bench_ab.json 2>&1; python - <<PY
import json
d=json.loads(open('gpurun_out/bench_ab.json').read().strip().splitlines()[-1])
print('xpf $x $w', 'ms/iter', round(d['ms_per_step'],3), 'G1', round(d['roofline']['launch_ms'],3), 'G2/3 avg', round(d['roofline']['other_gemm_launch_ms'],3), d['clocks']['sm_mhz'])
PY
done
done
done
