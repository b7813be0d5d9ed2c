#!/bin/bash
mkdir -p gpurun_out
python tools/tune_accept.py 1e-7 > gpurun_out/r2d_tune.log 2>&1; cat gpurun_out/r2d_tune.log
timeout 2400 python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/r2d_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2d_pytest.log; tail -4 gpurun_out/r2d_pytest.log
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/r2d_bench.json 2> gpurun_out/r2d_bench.err; echo "bench rc=$?"
timeout 900 ncu --set full --import-source on --clock-control none --kernel-name regex:mesh_level --launch-skip 4 --launch-count 1 \
    -o gpurun_out/r2d_mesh -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-state-points > /dev/null 2> gpurun_out/r2d_ncu_mesh.err
