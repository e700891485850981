#!/bin/bash
# round 2, GPU call 1: the whole -m gpu suite (new headline / restarts / patched-reference tests included), then the default bench
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r2_run1_gpus.txt; nproc >> gpurun_out/r2_run1_gpus.txt
timeout 2400 python -m pytest tests -m gpu -q --durations=25 -p no:cacheprovider > gpurun_out/r2_pytest1.log 2>&1
echo "pytest exit $?" >> gpurun_out/r2_pytest1.log
tail -60 gpurun_out/r2_pytest1.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r2_bench_c4_n1.json 2> gpurun_out/r2_bench_c4_n1.err
echo "bench exit $?"; tail -c 3000 gpurun_out/r2_bench_c4_n1.json; tail -5 gpurun_out/r2_bench_c4_n1.err
