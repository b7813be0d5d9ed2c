#!/bin/bash
# Turns the scratch captures of tools/wrap_profiles_r2.sh (gpurun_out/<tag>_*) into the tracked evidence under profiles/:
# bench lines, test log, launch list, FP64 op counters, and per kernel the ncu raw page (CSV), a digest and the hottest source lines.
tag=${1:-r2}
cd "$(dirname "$0")/.."
mkdir -p profiles
for f in bench_n1.json bench_reference_arm.json pytest_gpu.log smoke.log launches.csv fp64_ops.csv fp64_ops_bench.json eos_fp64_ops.csv quantum_fp64_ops.csv quantum_ops.log; do
  [ -s gpurun_out/${tag}_$f ] && cp gpurun_out/${tag}_$f profiles/${tag}_$f
done
cp gpurun_out/${tag}_ncu_*_raw.csv gpurun_out/${tag}_ncu_*_digest.txt profiles/ 2>/dev/null
ls -la profiles | grep ${tag}_
