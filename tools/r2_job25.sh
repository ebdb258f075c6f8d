#!/bin/bash
# after the last change to the map builder: parity, the full-size stress test, the default bench line
mkdir -p gpurun_out/j25
timeout 400 python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py -x -q > gpurun_out/j25/pytest_parity.log 2>&1; echo "parity rc $?"; tail -2 gpurun_out/j25/pytest_parity.log
timeout 300 python -m pytest tests/test_gpu_fullsize.py -x -q -k "stress_frame_tiles or textured_mesh" > gpurun_out/j25/pytest_full.log 2>&1; echo "fullsize rc $?"; tail -2 gpurun_out/j25/pytest_full.log
python bench.py --steps 20 --warmup 5 --no-cpu > gpurun_out/j25/bench.json 2> gpurun_out/j25/bench.err; echo "bench rc $?"; cut -c1-260 gpurun_out/j25/bench.json
