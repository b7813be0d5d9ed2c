#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_quantum.py -m gpu -q -s -p no:cacheprovider > gpurun_out/quantum_pytest.log 2>&1
echo "pytest rc=$?"; grep -v "^$" gpurun_out/quantum_pytest.log | tail -40 | cut -c1-260
