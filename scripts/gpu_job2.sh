set -x
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "raw_gemm" > gpurun_out/pytest_raw.log 2>&1; echo "raw rc=$?"; tail -25 gpurun_out/pytest_raw.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/bench_pair.json 2> gpurun_out/bench_pair.err; echo "bench rc=$?"; cat gpurun_out/bench_pair.json; tail -5 gpurun_out/bench_pair.err
NMF_UMMA_PAIR=0 timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e > gpurun_out/bench_nopair_hints.json 2>&1; cat gpurun_out/bench_nopair_hints.json
