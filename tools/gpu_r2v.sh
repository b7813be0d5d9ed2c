#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_regimes.py tests/test_gpu_chi_omega.py tests/test_gpu_generic.py tests/test_gpu_qmie.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -4 | cut -c1-220
bash tools/gpu_quick.sh base
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/r2v_launches.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-state-points --no-extra-configs > /dev/null 2>&1
python tools/launch_shares.py gpurun_out/r2v_launches.csv
