#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_quantum.py tests/test_gpu_generic.py tests/test_gpu_qmie.py -m gpu -q -s -p no:cacheprovider > gpurun_out/r2w_pytest.log 2>&1
echo "pytest rc=$?"; grep "max\|err\|passed\|failed" gpurun_out/r2w_pytest.log | tail -14 | cut -c1-250
