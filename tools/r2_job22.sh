#!/bin/bash
# 8-GPU box, after the occluder maps: the multi-GPU tests, the scaling lines (N = 2, 4, 8; N = 1 is tools/r2_job19.sh's
# line on the same hardware) and the single-process multi-GPU ABI
mkdir -p gpurun_out/j22
timeout 600 python -m pytest tests/test_multigpu_abi.py tests/test_multigpu_nccl.py -m gpu -x -q > gpurun_out/j22/pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/j22/pytest.log
run() { # workload N extra-args
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $2 --master-addr 127.0.0.1 --master-port 296$2$2 bench.py --gpus $2 --workload $1 --steps 20 --warmup 5 --no-cpu $3 > gpurun_out/j22/r2_scale_${1}_${2}gpu.json 2> gpurun_out/j22/r2_scale_${1}_${2}gpu.err
  echo "$1 N=$2 rc=$?"; cut -c1-200 gpurun_out/j22/r2_scale_${1}_${2}gpu.json
}
for n in 8 4 2; do run stress_1M_4k $n; done
run stress_1M_8k 8 --no-e2e; run test7_4k 8 --no-e2e; run house_4k 8 --no-e2e
timeout 300 python tools/multi_probe.py 8 > gpurun_out/j22/r2_multi_probe_8gpu.json 2> gpurun_out/j22/r2_multi_probe_8gpu.err
timeout 300 python tools/multi_probe.py 8 7680 4320 > gpurun_out/j22/r2_multi_probe_8gpu_8k.json 2>> gpurun_out/j22/r2_multi_probe_8gpu.err
tail -c 400 gpurun_out/j22/r2_multi_probe_8gpu.json
