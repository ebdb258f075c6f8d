#!/bin/bash
# Round-end scaling table on an 8-GPU box: N = 1, 2, 4, 8 (torchrun, one rank per GPU, NCCL allgather) + the
# single-process multi-GPU ABI.  stress_1M at 4K and 8K, test7 and house at 4K.
mkdir -p gpurun_out
run() { # workload N extra-args
  if [ $2 = 1 ]; then python bench.py --gpus 1 --workload $1 --steps 20 --warmup 5 --no-cpu $3 > gpurun_out/r2_scale_${1}_1gpu.json 2> gpurun_out/r2_scale_${1}_1gpu.err
  else python -m torch.distributed.run --nnodes=1 --nproc-per-node $2 --master-addr 127.0.0.1 --master-port 296$2$2 bench.py --gpus $2 --workload $1 --steps 20 --warmup 5 --no-cpu $3 > gpurun_out/r2_scale_${1}_${2}gpu.json 2> gpurun_out/r2_scale_${1}_${2}gpu.err; fi
  echo "$1 N=$2 rc=$?"
}
for n in 1 2 4 8; do run stress_1M_4k $n; done
# (N = 1 of these three: tools/r2_measure.sh, same hardware)
run test7_4k 8 --no-e2e; run house_4k 8 --no-e2e; run stress_1M_8k 8 --no-e2e
python tools/multi_probe.py 8 > gpurun_out/r2_multi_probe_8gpu.json 2> gpurun_out/r2_multi_probe_8gpu.err
python tools/multi_probe.py 8 7680 4320 > gpurun_out/r2_multi_probe_8gpu_8k.json 2>> gpurun_out/r2_multi_probe_8gpu.err
