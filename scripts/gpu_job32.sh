timeout 300 python tests/multi_gpu_update_div.py 2 > gpurun_out/mgpu_update_div.log 2>&1; echo "mgpu rc=$?"; tail -5 gpurun_out/mgpu_update_div.log | cut -c1-600
