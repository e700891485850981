# round evidence on one B200: GPU test suite, one bench line per BASELINE workload, the C4 bench line (the driver's
# command), the ncu launch list of the same command and one full capture of the two fused half-iteration kernels
# (each ncu pass only after the plain command has exited 0)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log | cut -c1-300
for w in c1 c2 c3 c5k64 c5 c5k512; do
timeout 600 python bench.py --workload $w --steps 50 --warmup 3 --no-cpu --no-e2e > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/bench_$w.json').read().strip().splitlines()[-1])
    print('$w', 'ms/iter', round(d['ms_per_step'],4), 'eff TF/s', round(d['effective_tflops'],1), 'fused', d['roofline']['fused_steps'], 'launches/iter', d['gpu_launches']//50, 'phases us', [round(x/3*1e6,1) for x in d['roofline']['phase_seconds_3_iterations'][:6]], (d['clocks'] or {}).get('sm_mhz'))
except Exception as e:
    print('$w bench failed', e); print(open('gpurun_out/bench_$w.err').read()[-600:])
PY
done
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_c4_n1.json 2> gpurun_out/bench_c4_n1.err; echo "bench rc=$?"; cat gpurun_out/bench_c4_n1.json | cut -c1-2500
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_arm.json 2> gpurun_out/bench_ref_arm.err; echo "reference arm rc=$?"; cut -c1-600 gpurun_out/bench_ref_arm.json
timeout 600 python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/plain.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1_launches_c4.csv python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_fused -s 6 -c 2 -o gpurun_out/r1_fused_c4 -f python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_full_c4.log 2>&1; echo "ncu full c4 rc=$?"
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
