#!/bin/bash
mkdir -p gpurun_out
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2t_bench.json 2> gpurun_out/r2t_bench.err; echo "bench rc=$?"; tail -5 gpurun_out/r2t_bench.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r2t_bench.json").read().strip().splitlines()[-1])
print("headline ms", d["ms_per_step"], d["value"])
print(json.dumps(d.get("eos_inputs"), indent=1)[:2500])
print("five", d["five_component"]["value"], d["five_component"]["ms_per_step"])
PY
