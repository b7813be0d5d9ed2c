#!/bin/bash
# usage: gpu_scale.sh N   (run under gpurun --gpus N)
N=$1
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench_n$N.json 2> gpurun_out/r2_bench_n$N.err
echo "bench n$N rc=$?"; tail -3 gpurun_out/r2_bench_n$N.err
python - $N <<'PY'
import json, sys
d = json.loads(open("gpurun_out/r2_bench_n%s.json" % sys.argv[1]).read().strip().splitlines()[-1])
print("N", d["n_gpus"], "ms", d["ms_per_step"], "value", d["value"], "e2e", d["e2e"]["value"], "lb", d.get("load_balance", {}).get("max_over_mean"))
for k in ("state_points", "five_component", "qmie_sweep"):
    if k in d: print(k, round(d[k]["value"]), d[k]["ms_per_step"])
s = d.get("strong_scaling", {})
for k, v in s.items(): print("strong", k, {kk: vv for kk, vv in v.items() if kk in ("value", "ms_per_step", "unit")})
PY
