#!/bin/bash
# Round 2, job 11: spawn kernel with all loads issued at once; separate trace / spawn times.
P="python tools/perf_probe.py --w 3840 --h 2160 --reps 4 --no-instrument"
$P > gpurun_out/r2_j11_profile.json 2> gpurun_out/err_j11.log
$P --nop > gpurun_out/r2_j11_nop.json 2>> gpurun_out/err_j11.log
$P --nop --world 8 --rank 3 > gpurun_out/r2_j11_shard8.json 2>> gpurun_out/err_j11.log
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "generated or example_scenes or secondary" > gpurun_out/r2_j11_pytest.log 2>&1; tail -2 gpurun_out/r2_j11_pytest.log
