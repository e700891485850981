export NMF_UMMA_FUSE=2
timeout 600 python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/plain_fused.log 2>&1 || exit 1
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_fused_w -s 3 -c 1 -o gpurun_out/fused_w -f python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_fused.log 2>&1; echo "ncu rc=$?"
