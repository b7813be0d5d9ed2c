#!/bin/bash
# quantum engine: parity suite + the bench leg's timing (three repeats).  usage: gpu_quantum_time.sh [tag]
tag=${1:-q}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_quantum.py -m gpu -q -p no:cacheprovider > gpurun_out/${tag}_pytest.log 2>&1
echo "pytest rc=$?"; tail -3 gpurun_out/${tag}_pytest.log
timeout 600 python - <<'PY'
import bench
for _ in range(3):
    r = bench.run_quantum_leg(0)
    print("quantum leg:", round(r["value"], 1), r["unit"], round(r["seconds"], 3), "s", {k: v for k, v in r.items() if "parity" in k or "err" in k})
PY
