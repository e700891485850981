#!/bin/bash
# round 2, GPU call 11: K = 128 (C3) with the correction passes in BF16 (NMF_CORR_BF16=3) against the default; C3 as
# specified (convergence run); the fused kernels on one rank's shard of the 8-GPU run (c4s8); C4 line with the new keys
mkdir -p gpurun_out
show() {
python - "$1" "$2" <<'PY'
import json, sys
tag, path = sys.argv[1], sys.argv[2]
try:
    d = json.loads(open(path).read().strip().splitlines()[-1])
    r = d['roofline']
    print(tag, 'ms/iter', round(d['ms_per_step'], 4), 'sustained', round((d.get('sustained') or {}).get('ms_per_step', 0), 4), 'eff TF/s', round(d['effective_tflops'], 1),
          'frac', round(r['frac'], 3), 'frac_3xtf32', round(r.get('frac_3xtf32', 0), 3), 'corr', r.get('corr_bf16'),
          'phases us', [round(x / 3 * 1e6, 1) for x in r['phase_seconds_3_iterations'][:8]], (d['clocks'] or {}).get('sm_mhz'))
except Exception as e:
    print(tag, 'bench failed', e); print(open(path.replace('.json', '.err')).read()[-800:])
PY
}
for rep in 1 2; do
for c in 1 3; do
  NMF_CORR_BF16=$c timeout 600 python bench.py --workload c3 --steps 100 --warmup 5 --no-cpu --no-e2e --no-refcuda > gpurun_out/r2_bench_c3_corr$c.json 2> gpurun_out/r2_bench_c3_corr$c.err
  show "c3 corr=$c" gpurun_out/r2_bench_c3_corr$c.json
done
done
for c in 1 3; do
  MARGIN_CHUNKS=4 NMF_CORR_BF16=$c timeout 600 python scripts/parity_margin.py > gpurun_out/r2_margin_c3_corr$c.jsonl 2> gpurun_out/r2_margin_c3_corr$c.err; cat gpurun_out/r2_margin_c3_corr$c.jsonl | cut -c1-500; tail -2 gpurun_out/r2_margin_c3_corr$c.err
done
timeout 600 python bench.py --workload c4 --steps 20 --warmup 3 > gpurun_out/r2_bench_c4_n1.json 2> gpurun_out/r2_bench_c4_n1.err; show c4 gpurun_out/r2_bench_c4_n1.json
timeout 600 python bench.py --workload c4s8 --steps 50 --warmup 3 --no-cpu --no-e2e --no-refcuda > gpurun_out/r2_bench_c4s8.json 2> gpurun_out/r2_bench_c4s8.err; show c4s8 gpurun_out/r2_bench_c4s8.json
timeout 900 python scripts/run_c3_convergence.py > gpurun_out/r2_c3_convergence.json 2> gpurun_out/r2_c3_convergence.err; cut -c1-700 gpurun_out/r2_c3_convergence.json; tail -2 gpurun_out/r2_c3_convergence.err
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_fused -s 6 -c 2 -o gpurun_out/r2_fused_c4s8 -f python bench.py --workload c4s8 --steps 2 --warmup 3 --no-e2e --no-cpu --no-refcuda --no-sustained > gpurun_out/ncu_full_c4s8.log 2>&1; echo "ncu c4s8 rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r2_launches_c4s8.csv python bench.py --workload c4s8 --steps 2 --warmup 3 --no-e2e --no-cpu --no-refcuda --no-sustained > /dev/null 2>&1; echo "ncu launches c4s8 rc=$?"
