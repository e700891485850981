#!/bin/bash
# round 2, GPU call 25 (2 GPUs): the K > 256 two-kernel path with BF16 corrections on a sharded run
mkdir -p gpurun_out
timeout 600 python tests/multi_gpu_update_div.py 2 512 0 > gpurun_out/r2_update_div_num_gpus2_k512.log 2>&1; grep PARITY gpurun_out/r2_update_div_num_gpus2_k512.log | cut -c1-500; tail -2 gpurun_out/r2_update_div_num_gpus2_k512.log | cut -c1-300
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29811 bench.py --gpus 2 --workload c5wk512 --steps 20 --warmup 3 --no-e2e > gpurun_out/r2_bench_c5wk512_n2.json 2> gpurun_out/r2_bench_c5wk512_n2.err
python - <<'PY'
import json
try:
    d=json.loads([l for l in open('gpurun_out/r2_bench_c5wk512_n2.json').read().strip().splitlines() if l.startswith('{')][-1])
    print('c5wk512 n=2', 'ms/iter', round(d['ms_per_step'],4), 'sustained', round(d['sustained']['ms_per_step'],4), 'corr', d['roofline'].get('corr_bf16'), 'parity', d.get('parity_vs_1gpu'))
except Exception as e:
    print('bench failed', e); print(open('gpurun_out/r2_bench_c5wk512_n2.err').read()[-1500:])
PY
