#!/bin/bash
# first run of the one-launch closest-hit kernel: parity subset, then A/B against the per-generation launches
mkdir -p gpurun_out/j15
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -k "generated or fuzz or secondary or graph or subframe or sharding or example or deterministic" > gpurun_out/j15/pytest.log 2>&1
echo "pytest rc $?"; tail -15 gpurun_out/j15/pytest.log
export RT_B200_LIB=$PWD/simple_raytracer_b200/lib/librt_b200_lab.so
for one in 0 1; do
  for cfg in "--world 8 --rank 0" "--world 4 --rank 0" ""; do
    echo "== RT_ONE_LAUNCH=$one $cfg"
    RT_ONE_LAUNCH=$one timeout 300 python tools/perf_probe.py --w 3840 --h 2160 $cfg --reps 8 --nop --no-instrument 2>&1 | cut -c1-200
  done
done > gpurun_out/j15/ab.log 2>&1
cat gpurun_out/j15/ab.log
