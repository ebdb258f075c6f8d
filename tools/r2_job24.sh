#!/bin/bash
# last word: the whole GPU suite and the default bench line on the final tree
mkdir -p gpurun_out/j24
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/j24/smoke.log 2>&1; echo "smoke rc $?"
python bench.py --steps 20 --warmup 5 > gpurun_out/j24/r2_bench_1gpu_stress_1M_4k.json 2> gpurun_out/j24/bench.err; echo "bench rc $?"
python -m pytest tests -m gpu -x -q > gpurun_out/j24/pytest.log 2>&1; echo "pytest rc $?"; tail -4 gpurun_out/j24/pytest.log
