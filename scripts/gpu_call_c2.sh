mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log | cut -c1-300
bash scripts/ref_cuda_c2.sh 2>&1 | tail -40
for f in 0 2; do
NMF_UMMA_FUSE=$f timeout 300 python bench.py --workload c2 --steps 200 --warmup 3 --no-cpu --no-e2e > gpurun_out/bench_c2_f$f.json 2> gpurun_out/bench_c2_f$f.err; python - <<PY
import json
d=json.loads(open('gpurun_out/bench_c2_f$f.json').read().strip().splitlines()[-1])
print('c2 fuse=$f', 'ms/iter', round(d['ms_per_step'],4), 'launches', d['gpu_launches'], 'phases us', [round(x/3*1e6,1) for x in d['roofline']['phase_seconds_3_iterations']])
PY
done
NMF_UMMA_FUSE=0 timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -s 60 -c 64 --csv --log-file gpurun_out/ours_c2_launches.csv python bench.py --workload c2 --steps 20 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_c2.log 2>&1; echo "ncu ours rc=$?"
python - <<PY
import csv,collections
rows=list(csv.DictReader(l for l in open('gpurun_out/ours_c2_launches.csv') if l.startswith('"')))
agg=collections.OrderedDict()
for r in rows:
    k=r['Kernel Name'][:90]; v=float(r['Metric Value'])/1e3; agg.setdefault(k,[]).append(v)
for k,v in agg.items(): print(f"{k:92s} n={len(v):3d} avg={sum(v)/len(v):8.2f} us")
PY
