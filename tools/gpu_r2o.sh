#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_regimes.py tests/test_gpu_chi_omega.py -m gpu -x -q -p no:cacheprovider > gpurun_out/r2o_pytest.log 2>&1
echo "pytest rc=$?"; tail -3 gpurun_out/r2o_pytest.log
bash tools/gpu_quick.sh base
KGCUDA_LIB=$PWD/_variants/libkgcuda_mesh3.so bash tools/gpu_quick.sh mesh3
