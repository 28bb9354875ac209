#!/bin/bash
# GPU parity tests, then the bench line with per-phase times.
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python bench.py --steps 5 --warmup 3 2>&1 | tail -1 > gpurun_out/bench_last.json
python - <<'PY'
import json
d = json.load(open("gpurun_out/bench_last.json"))
print("ms/step", d["ms_per_step"], "e2e ms", d["e2e"]["ms_per_step"], {k: round(v["ms_per_launch"], 3) for k, v in d["phases"].items()})
PY
