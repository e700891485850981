# the other BASELINE workloads on one GPU, fused half-iterations off / default (same box), one line each
for w in ${WLS:-c3 c5 c5k64 c5k512}; do
for f in 0 1; do
NMF_UMMA_FUSE=$f timeout 600 python bench.py --workload $w --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/bench_${w}_f$f.json 2> gpurun_out/bench_${w}_f$f.err; python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/bench_${w}_f$f.json').read().strip().splitlines()[-1])
    r=d['roofline']
    print('$w fuse=$f', 'ms/iter', round(d['ms_per_step'],4), 'eff TF/s', round(d['effective_tflops'],1), 'fused_steps', r['fused_steps'], 'launch_ms', round(r['launch_ms'],3), 'phases ms', [round(x/3*1e3,3) for x in r['phase_seconds_3_iterations'][:6]], d['clocks']['sm_mhz'])
except Exception as e:
    print('$w bench failed', e); print(open('gpurun_out/bench_${w}_f$f.err').read()[-800:])
PY
done
done
