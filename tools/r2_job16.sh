#!/bin/bash
# one-launch closest-hit kernel: parity subset, A/B against the per-generation launches, one full ncu capture
mkdir -p gpurun_out/j16
RT_B200_LIB=$PWD/simple_raytracer_b200/lib/librt_b200_lab.so RT_ONE_LAUNCH=1 timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -k "generated or fuzz or secondary or graph or subframe or sharding or example or deterministic" > gpurun_out/j16/pytest.log 2>&1
echo "pytest rc $?"; tail -3 gpurun_out/j16/pytest.log
export RT_B200_LIB=$PWD/simple_raytracer_b200/lib/librt_b200_lab.so
for one in 0 1; do
  for cfg in "--world 8 --rank 0" ""; do
    echo "== RT_ONE_LAUNCH=$one $cfg"
    RT_ONE_LAUNCH=$one timeout 300 python tools/perf_probe.py --w 3840 --h 2160 $cfg --reps 8 --nop --no-instrument 2>&1 | cut -c1-200
  done
done > gpurun_out/j16/ab.log 2>&1
cat gpurun_out/j16/ab.log
P="python tools/perf_probe.py --w 3840 --h 2160 --world 4 --rank 0 --reps 2 --nop --no-instrument"
RT_ONE_LAUNCH=1 timeout 300 $P > gpurun_out/j16/plain.log 2>&1 &&
RT_ONE_LAUNCH=1 timeout 900 ncu --set full --clock-control none --import-source on -f -k regex:bvh_trace_all_kernel -s 1 -c 1 -o gpurun_out/j16/trace_all $P > gpurun_out/j16/ncu.log 2>&1
ls -la gpurun_out/j16
