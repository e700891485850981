timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_gpu.log | cut -c1-300
timeout 900 python scripts/run_c3_convergence.py > gpurun_out/c3_convergence.json 2>&1; cat gpurun_out/c3_convergence.json | tail -3
