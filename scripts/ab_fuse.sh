timeout 300 python -m pytest tests/test_gpu_parity.py -q -x -k "fused_w" 2>&1 | tail -5
for f in ${FUSE_LIST:-0 2}; do
NMF_UMMA_FUSE=$f timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu --no-e2e > gpurun_out/fuse_$f.json 2> gpurun_out/fuse_$f.err; echo "rc=$?"; tail -2 gpurun_out/fuse_$f.err
python - <<PY
import json
d=json.loads(open('gpurun_out/fuse_$f.json').read().strip().splitlines()[-1])
print('fuse=$f ms/iter', round(d['ms_per_step'],3), 'launches', d['gpu_launches'], 'phases ms', [round(x/3*1e3,3) for x in d['roofline']['phase_seconds_3_iterations']], d['clocks'])
PY
done
