#!/bin/bash
# 1/2/4/8-GPU bench lines on ONE box (what the driver's scaling run does).
python bench.py --gpus 1 --no-cpu > gpurun_out/r1_bench_1gpu_scale.json 2> gpurun_out/scale.err
for n in 2 4 8; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) \
      bench.py --gpus $n > gpurun_out/r1_bench_${n}gpu.json 2>> gpurun_out/scale.err
done
for f in gpurun_out/r1_bench_1gpu_scale.json gpurun_out/r1_bench_2gpu.json gpurun_out/r1_bench_4gpu.json gpurun_out/r1_bench_8gpu.json; do
  python -c "
import json,sys
d=json.loads(open('$f').read().strip().splitlines()[-1])
print(d['n_gpus'], 'ms', round(d['ms_per_step'],3), 'Mrays/s', round(d['value'],1), 'e2e ms', round(d['e2e']['ms_per_step'],2), 'launches', d['gpu_launches'])"
done
