#!/bin/bash
# shim tests onward + launch list of the headline leg after the sync-free change
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_shim.py tests/test_gpu_table.py tests/test_gpu_transport.py tests/test_gpu_xsec.py tests/test_gpu_tl.py tests/test_gpu_sharding.py -m gpu -q -p no:cacheprovider > gpurun_out/r2l_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2l_pytest.log; tail -8 gpurun_out/r2l_pytest.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/r2l_launches.csv \
   python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-state-points --no-extra-configs > gpurun_out/r2l_ncu.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/r2l_ncu.log | head -c 600
