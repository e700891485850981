#!/bin/bash
# round 2, GPU call 4: full ncu captures (with source) of the K <= 128 fused kernels after the quotient moved to tensor memory
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_fused -s 6 -c 2 -o gpurun_out/r2_fused_c5k64_yt -f python bench.py --workload c5k64 --steps 2 --warmup 3 --no-e2e --no-cpu --no-refcuda --no-sustained > gpurun_out/ncu_full_c5k64_yt.log 2>&1; echo "ncu c5k64 rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_fused -s 6 -c 2 -o gpurun_out/r2_fused_c3_yt -f python bench.py --workload c3 --steps 2 --warmup 3 --no-e2e --no-cpu --no-refcuda --no-sustained > gpurun_out/ncu_full_c3_yt.log 2>&1; echo "ncu c3 rc=$?"
ls -la gpurun_out/*.ncu-rep
