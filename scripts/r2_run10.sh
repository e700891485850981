#!/bin/bash
# round 2, GPU call 10: BF16 correction passes on by default: the whole GPU suite (headline margins re-measured), C4 / C5 bench
# lines with the reference CUDA build, full ncu capture of the C4 kernels
mkdir -p gpurun_out; rm -f gpurun_out/headline_margins.jsonl gpurun_out/patched_reference_parity.jsonl
timeout 1800 python -m pytest tests -m gpu -q -x -p no:cacheprovider > gpurun_out/r2_pytest10.log 2>&1; tail -6 gpurun_out/r2_pytest10.log | cut -c1-300
cat gpurun_out/headline_margins.jsonl | cut -c1-300
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench_c4_n1.json 2> gpurun_out/r2_bench_c4_n1.err; echo "c4 rc=$?"; cut -c1-300 gpurun_out/r2_bench_c4_n1.json
timeout 900 python bench.py --workload c5 --steps 50 --warmup 3 --no-cpu > gpurun_out/r2_bench_c5.json 2> gpurun_out/r2_bench_c5.err; cut -c1-300 gpurun_out/r2_bench_c5.json
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_fused -s 6 -c 2 -o gpurun_out/r2_fused_c4_cb -f python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu --no-refcuda --no-sustained > gpurun_out/ncu_full_c4_cb.log 2>&1; echo "ncu c4 rc=$?"
