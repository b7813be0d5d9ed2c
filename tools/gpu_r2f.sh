#!/bin/bash
mkdir -p gpurun_out
timeout 1800 python -m pytest tests/test_gpu_transport.py tests/test_gpu_qmie.py tests/test_gpu_qmie_ret.py tests/test_gpu_zarate.py tests/test_gpu_generic.py tests/test_reference_goldens.py tests/test_gpu_shim.py -m gpu -q -p no:cacheprovider > gpurun_out/r2f_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2f_pytest.log; tail -4 gpurun_out/r2f_pytest.log
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/r2f_bench.json 2> gpurun_out/r2f_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/r2f_bench.err
