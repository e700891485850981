#!/bin/bash
# round 2, GPU call 29: L2 -> SM delivery against the size of a copy (are the 4 KB TMA boxes of the MN-major tiles what holds the
# kernels at 12 TB/s?)
mkdir -p gpurun_out
B=nmf_gpu_b200/build/l2_to_sm_bw
for a in "48 4 8 1" "48 4 16 1" "48 4 32 1" "48 4 16 2" "48 8 8 1" "48 8 16 1" "48 16 8 1" "48 32 4 1" "48 64 2 1" "48 2 32 1"; do timeout 120 $B $a; done | tee gpurun_out/r2_l2_to_sm_bw_sizes.jsonl
