#!/bin/bash
mkdir -p gpurun_out
python tools/profile_eos.py
M=smsp__sass_thread_inst_executed_op_dfma_pred_on.sum,smsp__sass_thread_inst_executed_op_dmul_pred_on.sum,smsp__sass_thread_inst_executed_op_dadd_pred_on.sum,gpu__time_duration.sum
ncu --metrics $M --clock-control none --kernel-name regex:"eos_" --csv --log-file gpurun_out/r2_eos_fp64_ops.csv python tools/profile_eos.py > /dev/null 2>&1
ncu --set full --import-source on --clock-control none --kernel-name regex:eos_hess --launch-skip 2 --launch-count 1 -f -o gpurun_out/r2_eos_hess python tools/profile_eos.py > /dev/null 2> gpurun_out/r2_ncu_eos.err
ncu -i gpurun_out/r2_eos_hess.ncu-rep --page raw --csv > gpurun_out/r2_ncu_eos_hess_raw.csv 2>/dev/null
{ python tools/ncu_digest.py gpurun_out/r2_eos_hess.ncu-rep 50; echo; python tools/ncu_lines.py gpurun_out/r2_eos_hess.ncu-rep 25; } > gpurun_out/r2_ncu_eos_hess_digest.txt 2>&1
rm -f gpurun_out/r2_eos_hess.ncu-rep
head -30 gpurun_out/r2_ncu_eos_hess_digest.txt | cut -c1-160
