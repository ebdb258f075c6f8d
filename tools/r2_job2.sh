#!/bin/bash
# Round 2, job 2: the no-alloc test alone (fresh process), the GPU suite, probes to compare with round 1.
python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r2_j2_noalloc_alone.log 2>&1; echo "alone rc=$?" >> gpurun_out/r2_j2_noalloc_alone.log
python -m pytest tests -m gpu -x -q > gpurun_out/r2_j2_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_j2_pytest.log
python tools/perf_probe.py --w 3840 --h 2160 --reps 2 > gpurun_out/r2_j2_probe_4k.json 2> gpurun_out/err3.log
python tools/perf_probe.py --w 3840 --h 2160 --reps 3 --nop --no-instrument > gpurun_out/r2_j2_probe_4k_nop.json 2>> gpurun_out/err3.log
python tools/perf_probe.py --w 3840 --h 2160 --reps 3 --nop --no-instrument --world 8 --rank 3 > gpurun_out/r2_j2_probe_4k_shard8_nop.json 2> gpurun_out/err4.log
python bench.py --workload test7_4k --steps 5 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2_j2_test7_4k.json 2> gpurun_out/err5.log
tail -3 gpurun_out/r2_j2_noalloc_alone.log; tail -5 gpurun_out/r2_j2_pytest.log
