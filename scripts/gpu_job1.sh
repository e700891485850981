set -x
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_c4_n1.json 2> gpurun_out/bench_c4_n1.err; echo "bench rc=$?"; cat gpurun_out/bench_c4_n1.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1_launches_c4.csv python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_umma -s 6 -c 3 -o gpurun_out/r1_umma_mid -f python bench.py --workload mid --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_full_mid.log 2>&1; echo "ncu full mid rc=$?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_umma -s 6 -c 3 -o gpurun_out/r1_umma_c4 -f python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_full_c4.log 2>&1; echo "ncu full c4 rc=$?"
ls -la gpurun_out
