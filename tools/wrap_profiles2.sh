set -x
python -m pytest tests/test_gpu_qmie.py -q 2>&1 | tail -2
python bench.py > gpurun_out/r1d_bench_n1.json 2> gpurun_out/r1d_bench_n1.err
ncu --set full --clock-control none --import-source on -k regex:theta_kernel --launch-skip 25 --launch-count 3 -f -o gpurun_out/r1d_theta python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extra-configs > gpurun_out/r1d_ncu2.log 2>&1
