#!/bin/bash
# quantum leg with the cross-section kernel compiled for 4 / 6 / 8 CTAs per SM (tools/build_variant.sh q6 -DKG_Q_MIN_CTAS=6 ...)
for v in "" q6 q8; do
lib=""; [ -n "$v" ] && lib=_variants/libkgcuda_$v.so
KGCUDA_LIB=$lib timeout 300 python - <<PY
import bench
r = [bench.run_quantum_leg(0) for _ in range(2)][-1]
print("variant '$v':", round(r["value"], 1), r["unit"], round(r["seconds"], 3), "s", r["parity"]["max_rel_err"])
PY
done
