#!/bin/bash
# source-level ncu captures of the three hot kernels at the largest pass (5th launch) of the headline run
mkdir -p gpurun_out
for k in theta_setup theta_kernel mesh_level; do
timeout 900 ncu --set full --import-source on --clock-control none --kernel-name regex:$k --launch-skip 4 --launch-count 1 \
    -o gpurun_out/r2m_$k -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-state-points --no-extra-configs > /dev/null 2> gpurun_out/r2m_ncu_$k.err
echo "$k rc=$?"
done
ls -la gpurun_out/r2m*.ncu-rep
