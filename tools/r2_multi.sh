#!/bin/bash
# Round 2: the multi-GPU paths on a box with N >= 2 GPUs (gpurun --gpus N): ABI tests, torchrun NCCL test, bench lines.
N=${1:-2}
python -m pytest tests/test_multigpu_abi.py tests/test_multigpu_nccl.py -m gpu -x -q > gpurun_out/r2_multi_pytest_$N.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_multi_pytest_$N.log
tail -4 gpurun_out/r2_multi_pytest_$N.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r2_multi_bench_${N}gpu.json 2> gpurun_out/r2_multi_bench_${N}gpu.err
tail -c 600 gpurun_out/r2_multi_bench_${N}gpu.json
python tools/multi_probe.py $N > gpurun_out/r2_multi_probe_${N}gpu.json 2> gpurun_out/r2_multi_probe_${N}gpu.err; tail -c 900 gpurun_out/r2_multi_probe_${N}gpu.json
