#!/bin/bash
# headline leg only: step time, stage split, parity block.  usage: gpu_quick.sh [tag]   (KGCUDA_LIB selects a variant build)
tag=${1:-quick}
mkdir -p gpurun_out
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-state-points --no-extra-configs > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err
echo "[$tag] bench rc=$?"; tail -3 gpurun_out/${tag}_bench.err
python - $tag <<'PY'
import json, sys
d = json.loads(open("gpurun_out/%s_bench.json" % sys.argv[1]).read().strip().splitlines()[-1])
r = d["roofline"]
print("ms/step", round(d["ms_per_step"], 3), "value", round(d["value"]), "e2e", round(d["e2e"]["value"]),
      "| theta+setup ms", round(r["kernel_ms_per_step"], 3), "mesh ms", round(r["mesh_ms_per_step"], 3), "| parity", d["parity"]["max_rel_err_per_regime"])
PY
