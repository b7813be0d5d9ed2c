#!/bin/bash
for a in 3 4 5 6 8; do
KG_Q_AHEAD=$a timeout 300 python - <<PY
import bench
r = [bench.run_quantum_leg(0) for _ in range(2)][-1]
print("AHEAD $a:", round(r["value"], 1), r["unit"], round(r["seconds"], 3), "s", r["parity"]["max_rel_err"])
PY
done
