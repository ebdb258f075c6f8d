#!/bin/bash
# final verification: smoke, the whole GPU suite, the default bench line and the reference arm
mkdir -p gpurun_out/j12
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/j12/smoke.log 2>&1
echo "smoke rc $?"
python -m pytest tests -m gpu -x -q > gpurun_out/j12/pytest.log 2>&1
echo "pytest rc $?"; tail -5 gpurun_out/j12/pytest.log
python bench.py --steps 20 --warmup 5 > gpurun_out/j12/bench_stress_1M_4k.json 2> gpurun_out/j12/bench.err
echo "bench rc $?"; cat gpurun_out/j12/bench_stress_1M_4k.json
python tools/perf_probe.py > gpurun_out/j12/perf_probe.log 2>&1; tail -30 gpurun_out/j12/perf_probe.log
