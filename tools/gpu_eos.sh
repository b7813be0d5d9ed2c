#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_eos.py -m gpu -q -s -p no:cacheprovider > gpurun_out/eos_pytest.log 2>&1
echo "pytest rc=$?"; grep -v "^$" gpurun_out/eos_pytest.log | tail -30 | cut -c1-250
