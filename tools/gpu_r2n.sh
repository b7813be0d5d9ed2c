#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_regimes.py tests/test_gpu_chi_omega.py tests/test_gpu_generic.py -m gpu -x -q -p no:cacheprovider > gpurun_out/r2n_pytest.log 2>&1
echo "pytest rc=$?"; tail -3 gpurun_out/r2n_pytest.log
bash tools/gpu_quick.sh base
KGCUDA_LIB=$PWD/_variants/libkgcuda_mesh3.so bash tools/gpu_quick.sh mesh3
KGCUDA_LIB=$PWD/_variants/libkgcuda_mesh2.so bash tools/gpu_quick.sh mesh2
for p in 3 4; do
timeout 600 ncu --set full --import-source on --clock-control none --kernel-name regex:mesh_level --launch-skip $p --launch-count 1 \
    -o gpurun_out/r2n_mesh_p$p -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-state-points --no-extra-configs > /dev/null 2> gpurun_out/r2n_ncu.err
done
