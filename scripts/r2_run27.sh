#!/bin/bash
# round 2, GPU call 27: round-end rehearsal with the final tree (pytest -m gpu, smoke, both bench arms) + one full capture of the
# three kernels of the K = 512 two-kernel path with the BF16 corrections
mkdir -p gpurun_out
timeout 1800 python -m pytest tests/ -x -q -m gpu -p no:cacheprovider > gpurun_out/r2_pytest_gpu.log 2>&1; tail -3 gpurun_out/r2_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 600 python bench.py --impl reference --gpus 1 --steps 2 --warmup 1 > gpurun_out/r2_bench_reference_arm.json 2>/dev/null; cut -c1-200 gpurun_out/r2_bench_reference_arm.json
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2_bench_c4_n1.json 2> gpurun_out/r2_bench_c4_n1.err; python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2_bench_c4_n1.json').read().strip().splitlines()[-1]); r=d['roofline']
print('c4 default', d['ms_per_step'], d['value'], d['e2e']['value'], r['frac'], r['frac_3xtf32'], d['clocks'], d['gpu_launches'], (d.get('reference_cuda') or {}).get('ms_per_step'))
PY
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_umma -s 12 -c 4 -o gpurun_out/r2_umma_c5k512 -f python bench.py --workload c5k512 --steps 2 --warmup 3 --no-e2e --no-cpu --no-refcuda --no-sustained > gpurun_out/ncu_full_c5k512.log 2>&1; echo "ncu c5k512 rc=$?"
