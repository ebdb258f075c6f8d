for im in 1 4 8 12 16 24; do echo "INNER_MIN=$im"; RT_INNER_MIN_LANES=$im timeout 100 python tools/perf_probe.py --reps 3 --no-instrument; done
for lm in 4 8 16; do echo "LEAF_MIN=$lm"; RT_LEAF_MIN_LANES=$lm timeout 100 python tools/perf_probe.py --reps 3 --no-instrument; done
for rf in 4 12 16; do echo "REFILL=$rf"; RT_REFILL_LANES=$rf timeout 100 python tools/perf_probe.py --reps 3 --no-instrument; done
