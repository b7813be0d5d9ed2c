#!/bin/bash
# sync-free pass loop: full GPU suite, then the bench line (no CPU legs)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/r2k_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2k_pytest.log; tail -15 gpurun_out/r2k_pytest.log
timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2k_bench.json 2> gpurun_out/r2k_bench.err
echo "bench rc=$?"; tail -3 gpurun_out/r2k_bench.err; head -c 3000 gpurun_out/r2k_bench.json
