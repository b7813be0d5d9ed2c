#!/bin/bash
mkdir -p gpurun_out
python tools/tune_accept.py 1e-9 1e-8 1e-7 1e-6 > gpurun_out/r2_tune_accept.log 2>&1; cat gpurun_out/r2_tune_accept.log
python tools/profile_tables.py > gpurun_out/r2_tables.log 2>&1; cat gpurun_out/r2_tables.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r2_tables_launches.csv python tools/profile_tables.py > /dev/null 2>&1
echo "ncu rc=$?"
