#!/bin/bash
# round 2, multi-GPU call: N = $1 GPUs of one box.  Multi-GPU pytest (2 GPUs), then C4 strong scaling and the C5 weak sweep at every
# power of two up to N, each line with its parity_vs_1gpu word.
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L | head -8
timeout 1500 python -m pytest tests/test_gpu_multi.py -m gpu -q -p no:cacheprovider > gpurun_out/r2_pytest_multi.log 2>&1; tail -6 gpurun_out/r2_pytest_multi.log
port=29700
for g in 2 4 8; do
  [ $g -le $N ] || continue
  for w in c4 c5w c5wk64 c5wk512; do
    port=$((port+1))
    extra=""
    if [ -n "$QUICK" ]; then                       # an 8-GPU box is charged 8x: e2e only on the C4 lines, the K = 64 / 512 sweeps not again at 2 GPUs
      [ $w = c4 ] || extra="--no-e2e"
      if [ $g -eq 2 ] && [ $w != c4 ] && [ $w != c5w ]; then continue; fi
    fi
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $g --master-addr 127.0.0.1 --master-port $port bench.py --gpus $g --workload $w --steps 20 --warmup 5 $extra > gpurun_out/r2_bench_${w}_n$g.json 2> gpurun_out/r2_bench_${w}_n$g.err
    python - <<PY
import json
try:
    d=json.loads([l for l in open('gpurun_out/r2_bench_${w}_n$g.json').read().strip().splitlines() if l.startswith('{')][-1])
    print('$w n=$g', 'ms/iter', round(d['ms_per_step'],4), 'it/s', round(d['value'],2), 'eff TF/s', round(d['effective_tflops'],1), 'sustained ms', round(d['sustained']['ms_per_step'],4), (d['sustained']['clocks'] or {}).get('sm_mhz'),
          'e2e', (d.get('e2e') or {}).get('value'), 'outside s', (d.get('e2e') or {}).get('outside_loop_seconds'), 'parity', d.get('parity_vs_1gpu'), 'comm ms/iter', round(d['roofline']['phase_seconds_3_iterations'][7]/3*1e3,3))
except Exception as e:
    print('$w n=$g bench failed', e); print(open('gpurun_out/r2_bench_${w}_n$g.err').read()[-1500:])
PY
  done
done
port=$((port+1))
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port tests/dist_gpu_parity.py > gpurun_out/r2_dist_parity_n$N.log 2>&1; grep PARITY gpurun_out/r2_dist_parity_n$N.log | cut -c1-600
timeout 600 python tests/multi_gpu_update_div.py $N 128 1 > gpurun_out/r2_update_div_num_gpus$N.log 2>&1; grep PARITY gpurun_out/r2_update_div_num_gpus$N.log | cut -c1-600
