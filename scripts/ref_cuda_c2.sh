# Times the reference CUDA program (oracle/_ref/nmf_ref_ndebug: the unmodified sources compiled -DNDEBUG, see
# oracle/Makefile) on BASELINE config C2 (513x20000, K=40, seed 1, 200 iterations) on this GPU, and our
# update_div on the same inputs.  Kernel times come from an ncu launch list (serialised, cold cache).
T=$(mktemp -d); mkdir -p $T/cuda
python - <<PY
import numpy as np, struct
np.random.seed(1)
X=np.random.rand(513,20000).astype(np.float32); W=np.random.rand(513,40).astype(np.float32); H=np.random.rand(40,20000).astype(np.float32)
for n,a in (("X",X),("W",W),("H",H)):
    with open("$T/%s.bin"%n,"wb") as f: f.write(struct.pack("ii",*a.shape)); f.write(a.tobytes())
PY
export LD_LIBRARY_PATH=/usr/local/cuda/lib64:$LD_LIBRARY_PATH NVIDIA_TF32_OVERRIDE=0
R=$PWD/oracle/_ref/nmf_ref_ndebug
OUT=$PWD/gpurun_out
( cd $T/cuda && $R > /dev/null && for i in 1 2 3; do /usr/bin/time -f "ref wall %e s" $R > /dev/null; done ) 2>&1 | grep "ref wall"
( cd $T/cuda && ncu --metrics gpu__time_duration.sum --clock-control none -s 40 -c 320 --csv --log-file $OUT/ref_c2_launches.csv $R > /dev/null 2>&1 ); echo "ncu ref rc=$?"
python - <<PY
import csv,collections
rows=list(csv.DictReader(l for l in open('gpurun_out/ref_c2_launches.csv') if l.startswith('"')))
agg=collections.OrderedDict(); tot=0
for r in rows:
    k=r['Kernel Name'][:70]; v=float(r['Metric Value'])/1e3; agg.setdefault(k,[]).append(v); tot+=v
for k,v in agg.items(): print(f"{k:72s} n={len(v):3d} avg={sum(v)/len(v):8.2f} us")
print('reference C2: launches',len(rows),'sum of kernel time',round(tot,1),'us ->', round(tot/len(rows)*16,1),'us per iteration of 16 graph nodes')
PY
python - <<PY
import numpy as np, time, sys, os
sys.path.insert(0,'.')
import nmf_gpu_b200.api as api
np.random.seed(1)
ld=lambda a: np.asfortranarray(a.reshape(-1).reshape(a.shape,order='F'))
X=ld(np.random.rand(513,20000).astype(np.float32)); W=ld(np.random.rand(513,40).astype(np.float32)); H=ld(np.random.rand(40,20000).astype(np.float32))
for fuse in ("0","2"):
    os.environ["NMF_UMMA_FUSE"]=fuse
    for mode in ("refcompat","exact"):
        for rep in range(3):
            Wd,Hd=W.copy(order='F'),H.copy(order='F')
            t0=time.perf_counter(); r=api.update_div(Wd,Hd,X,0.0,200,opt=api.options(semantics=mode)); dt=time.perf_counter()-t0
        print('ours fuse',fuse,mode,'200 iterations: device loop',round(r.device_ms,3),'ms ->',round(r.device_ms/200*1e3,1),'us/iteration; launches',r.gpu_launches,'; update_div wall',round(dt*1e3,2),'ms')
PY
