#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python bench.py > gpurun_out/r2y_bench.json 2> gpurun_out/r2y_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/r2y_bench.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r2y_bench.json").read().strip().splitlines()[-1])
print("headline", d["ms_per_step"], d["value"], d["e2e"]["value"], "cpu", d["cpu_baseline"]["value"])
print("quantum", d.get("quantum_omega"))
PY
timeout 600 python -m pytest tests/test_gpu_quantum.py -m gpu -q -p no:cacheprovider 2>&1 | tail -2
