set -x
N=${1:-2}
for f in 0 1; do   # two-kernel path / fused half-iterations (default)
NMF_UMMA_FUSE=$f timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$f tests/dist_gpu_parity.py > gpurun_out/dist_parity_n${N}_fuse$f.log 2>&1; echo "parity fuse=$f rc=$?"; grep PARITY gpurun_out/dist_parity_n${N}_fuse$f.log
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_c4_n$N.log 2>&1; echo "bench rc=$?"; grep '"metric"' gpurun_out/bench_c4_n$N.log
