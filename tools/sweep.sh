for lm in 1 2 4 8; do echo "LEAF_MAX=$lm"; RT_LEAF_MAX=$lm timeout 100 python tools/perf_probe.py --reps 2; done
for rf in 2 4 16 24; do echo "REFILL=$rf"; RT_REFILL_LANES=$rf timeout 100 python tools/perf_probe.py --reps 2 --no-instrument; done
for gm in 4 8 32; do echo "GRID_MULT=$gm"; RT_GRID_MULT=$gm timeout 100 python tools/perf_probe.py --reps 2 --no-instrument; done
