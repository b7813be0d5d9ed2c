#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/r2q_pytest.log 2>&1
echo "pytest rc=$?"; tail -3 gpurun_out/r2q_pytest.log
bash tools/gpu_quick.sh base
python - <<'PY'
import os, time, json, subprocess, sys
def leg(env):
    e = dict(os.environ, **env)
    out = subprocess.run([sys.executable, "bench.py", "--steps", "3", "--warmup", "3", "--no-cpu-baseline"], capture_output=True, text=True, env=e).stdout
    d = json.loads(out.strip().splitlines()[-1])
    for k in ("state_points", "five_component"):
        print(env, k, round(d[k]["value"]), "pts/s", round(d[k]["ms_per_step"], 3), "ms", d[k]["stage_ms"], "frac", d[k].get("roofline", {}).get("frac"))
    print(env, "headline", d["ms_per_step"], "qmie", d["qmie_sweep"]["value"], d["config0_latency"]["seconds_cold"])
    return d
a = leg({}); b = leg({"KG_TL_LANE_SEARCH": "1"})
json.dump(a, open("gpurun_out/r2q_bench.json", "w"))
PY
