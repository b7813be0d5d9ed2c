#!/bin/bash
# FP64 lane-instruction totals per kernel of the headline run (for bench.py's FP64_PER_* constants): ncu SASS op counters of every
# theta / mesh launch; the per-step work counters come from the bench line of the same command.
mkdir -p gpurun_out
M=smsp__sass_thread_inst_executed_op_dfma_pred_on.sum,smsp__sass_thread_inst_executed_op_dmul_pred_on.sum,smsp__sass_thread_inst_executed_op_dadd_pred_on.sum,gpu__time_duration.sum
timeout 900 ncu --metrics $M --clock-control none --kernel-name regex:"theta|mesh_level" -c 400 --csv --log-file gpurun_out/calib_fp64.csv \
   python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-state-points --no-extra-configs > gpurun_out/calib_bench.json 2> gpurun_out/calib.err
echo "rc=$?"; tail -c 300 gpurun_out/calib.err
