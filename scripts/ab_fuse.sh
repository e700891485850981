# A/B of the fused half-iterations at the bench workload, same box: "fuse:steps" pairs in FUSE_LIST
for fs in ${FUSE_LIST:-0:3 2:1 2:3}; do
f=${fs%%:*}; st=${fs##*:}
NMF_UMMA_FUSE=$f NMF_UMMA_FUSE_STEPS=$st timeout 600 python bench.py --workload ${WL:-c4} --steps ${STEPS:-5} --warmup 3 --no-cpu --no-e2e > gpurun_out/fuse_${f}_$st.json 2> gpurun_out/fuse_${f}_$st.err; echo "rc=$?"; tail -2 gpurun_out/fuse_${f}_$st.err
python - <<PY
import json
d=json.loads(open('gpurun_out/fuse_${f}_$st.json').read().strip().splitlines()[-1])
print('fuse=$f steps=$st ms/iter', round(d['ms_per_step'],3), 'launches', d['gpu_launches'], 'phases ms', [round(x/3*1e3,3) for x in d['roofline']['phase_seconds_3_iterations']], d['clocks'])
PY
done
