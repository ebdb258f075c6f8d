#!/bin/bash
# Round 2, job 9: the whole GPU suite on the final code, then one compute-sanitizer memcheck pass over a small case.
python -m pytest tests -m gpu -x -q > gpurun_out/r2_j9_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_j9_pytest.log
tail -4 gpurun_out/r2_j9_pytest.log
python tools/sanitize_case.py > gpurun_out/r2_j9_plain.log 2>&1 &&
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 python tools/sanitize_case.py > gpurun_out/r2_j9_memcheck.log 2>&1; echo "memcheck rc=$?"
tail -5 gpurun_out/r2_j9_memcheck.log
