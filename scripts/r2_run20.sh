#!/bin/bash
# round 2, GPU call 20: where the end-to-end call spends its time outside the device loop
mkdir -p gpurun_out
nvidia-smi topo -m 2>/dev/null | head -12
python - <<'PY'
import os
print('cpus', os.cpu_count(), 'affinity', len(os.sched_getaffinity(0)))
try:
    print(open('/sys/devices/system/node/online').read().strip(), [open(f'/sys/devices/system/node/node{i}/cpulist').read().strip() for i in range(4) if os.path.exists(f'/sys/devices/system/node/node{i}/cpulist')])
except Exception as e: print(e)
PY
timeout 600 python scripts/e2e_breakdown.py 8192 131072 256 50 | tee gpurun_out/r2_e2e_breakdown_c5.json
timeout 600 python scripts/e2e_breakdown.py 16384 262144 256 20 | tee gpurun_out/r2_e2e_breakdown_c4.json
timeout 600 python scripts/e2e_breakdown.py 513 20000 40 50 | tee gpurun_out/r2_e2e_breakdown_c2.json
