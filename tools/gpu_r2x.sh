#!/bin/bash
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/r2x_pytest.log 2>&1
echo "pytest rc=$?"; tail -4 gpurun_out/r2x_pytest.log
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2x_bench.json 2> gpurun_out/r2x_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/r2x_bench.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r2x_bench.json").read().strip().splitlines()[-1])
print("headline", d["ms_per_step"], d["value"], d["e2e"]["value"])
print("quantum", d.get("quantum_omega"))
print("eos", {k: d["eos_inputs"][k] for k in ("value", "ms_per_step", "kernel_ms")}, d["eos_inputs"]["roofline"]["frac"])
for k in ("state_points", "five_component", "qmie_sweep"): print(k, round(d[k]["value"]), d[k]["ms_per_step"])
PY
