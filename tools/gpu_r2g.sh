#!/bin/bash
mkdir -p gpurun_out
timeout 1800 python -m pytest tests/test_gpu_transport.py tests/test_gpu_qmie.py tests/test_gpu_qmie_ret.py tests/test_gpu_zarate.py tests/test_gpu_generic.py tests/test_reference_goldens.py -m gpu -q -p no:cacheprovider > gpurun_out/r2g_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2g_pytest.log; tail -4 gpurun_out/r2g_pytest.log
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/r2g_bench.json 2> gpurun_out/r2g_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/r2g_bench.err
timeout 600 ncu --set full --import-source on --clock-control none --kernel-name regex:state_kernel --launch-skip 2 --launch-count 1 -o gpurun_out/r2g_state -f python tools/profile_state.py > /dev/null 2> gpurun_out/r2g_ncu_state.err
