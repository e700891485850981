# one full ncu capture of a fused half-iteration kernel: WL (workload), STEPS_MASK (1 = W step, 2 = H step)
export NMF_UMMA_FUSE=2 NMF_UMMA_FUSE_STEPS=${STEPS_MASK:-1}
timeout 600 python bench.py --workload ${WL:-c4} --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/plain_fused.log 2>&1 || exit 1
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_fused -s 3 -c 1 -o gpurun_out/fused_${WL:-c4} -f python bench.py --workload ${WL:-c4} --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_fused.log 2>&1; echo "ncu rc=$?"
