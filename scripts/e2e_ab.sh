for f in 0 1; do
NMF_UMMA_FUSE=$f timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/e2e_$f.json 2> gpurun_out/e2e_$f.err; python - <<PY
import json
d=json.loads(open('gpurun_out/e2e_$f.json').read().strip().splitlines()[-1])
print('fuse=$f ms/iter', round(d['ms_per_step'],3), 'e2e', d['e2e'])
PY
done
