timeout 900 python bench.py --steps 200 --warmup 5 --no-cpu > gpurun_out/bench_c4_200.json 2> gpurun_out/bench_c4_200.err; python - <<PY
import json
d=json.loads(open('gpurun_out/bench_c4_200.json').read().strip().splitlines()[-1])
print('c4 200 steps', 'ms/iter', round(d['ms_per_step'],3), 'eff TF/s', round(d['effective_tflops'],1), 'e2e it/s', round(d['e2e']['value'],2), d['clocks'])
PY
python - <<PY
import sys, numpy as np
sys.path.insert(0,'.')
import nmf_gpu_b200.api as api
with api.Session(16384, 262144, 256) as s:
    s.fill_uniform(4)
    c=[s.cost()]
    for i in range(4):
        r=s.run(0.0, 50); c.append(s.cost())
    W,H=s.download()
print('KL every 50 iterations:', ['%.6e'%x for x in c], 'monotone', all(c[i+1]<=c[i] for i in range(len(c)-1)), 'finite', bool(np.isfinite(W).all() and np.isfinite(H).all()), 'min W', float(W.min()), 'min H', float(H.min()))
PY
make -C examples 2>&1 | tail -2; ls -la examples/nmf && (T=$(mktemp -d); mkdir $T/cuda; python - <<PY
import numpy as np, struct
np.random.seed(0)
X=np.random.rand(4096,350).astype(np.float32); W=np.random.rand(4096,128).astype(np.float32); H=np.random.rand(128,350).astype(np.float32)
for n,a in (("X",X),("W",W),("H",H)):
    with open("$T/%s.bin"%n,"wb") as f: f.write(struct.pack("ii",*a.shape)); f.write(a.tobytes())
PY
cd $T/cuda && NMF_SEMANTICS=refcompat $OLDPWD/examples/nmf; ls -la $T; cd $OLDPWD; python examples/compare_output.py $T/Wout.bin /root/reference/Wtest.bin 2>&1 | tail -2)
