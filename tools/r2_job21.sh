#!/bin/bash
# occluder maps, cheaper build: parity subset, then bench lines (with the e2e split) of the scenes that have maps
mkdir -p gpurun_out/j21
timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_multigpu_abi.py -x -q > gpurun_out/j21/pytest.log 2>&1
echo "pytest rc $?"; tail -4 gpurun_out/j21/pytest.log
for w in stress_1M_4k mesh_earth_1080p; do
  python bench.py --workload $w --steps 20 --warmup 5 --no-cpu > gpurun_out/j21/bench_$w.json 2> gpurun_out/j21/bench_$w.err
  echo "$w rc=$?"
  python - <<PY
import json
d=json.load(open('gpurun_out/j21/bench_$w.json')); k=d['roofline']['kernels']
print(d['ms_per_step'], {n:round(v['ms_per_frame'],3) for n,v in k.items()}, 'e2e', d['e2e']['ms_per_step'], d['e2e']['split_ms'], d['frame_check']['result'])
PY
done
python tools/perf_probe.py --w 3840 --h 2160 --world 8 --rank 0 --reps 8 --nop --no-instrument | cut -c1-100
