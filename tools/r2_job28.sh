#!/bin/bash
# the scanline map builder: parity subset, then the default bench line with its end-to-end split
timeout 70 python -m pytest tests/test_gpu_parity.py -x -q -k "occluder or fuzz or secondary" 2>&1 | tail -2
timeout 90 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/bench_final.json 2>/dev/null
python - <<'PY'
import json
d = json.load(open("gpurun_out/bench_final.json"))
print(d["ms_per_step"], d["e2e"]["ms_per_step"], d["e2e"]["split_ms"], d["frame_check"]["result"])
PY
