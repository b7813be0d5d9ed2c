#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/r2s_pytest.log 2>&1
echo "pytest rc=$?"; tail -3 gpurun_out/r2s_pytest.log
python tools/profile_tables.py 2>&1 | tail -2
timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2s_bench.json 2> gpurun_out/r2s_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r2s_bench.json").read().strip().splitlines()[-1])
print("headline ms", d["ms_per_step"], d["value"], d["e2e"]["value"], d["parity"]["max_rel_err"])
for k in ("state_points", "five_component", "qmie_sweep"):
    print(k, round(d[k]["value"]), d[k]["ms_per_step"], d[k].get("stage_ms"))
print(d["config0_latency"]["seconds_cold"])
PY
