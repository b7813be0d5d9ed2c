#!/bin/bash
# round 2, GPU call A: full GPU test suite, both bench arms, launch list of one full-size step
mkdir -p gpurun_out
{ nvidia-smi -L; nproc; lscpu | grep "Model name"; } > gpurun_out/r2a_env.txt 2>&1
timeout 2400 python -m pytest tests -m gpu -q -s -p no:cacheprovider > gpurun_out/r2a_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2a_pytest.log
tail -5 gpurun_out/r2a_pytest.log
timeout 900 python bench.py > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err
echo "bench rc=$?"; tail -3 gpurun_out/r2a_bench.err
timeout 600 python bench.py --impl reference --steps 1 --warmup 1 > gpurun_out/r2a_bench_ref.json 2> gpurun_out/r2a_bench_ref.err
echo "ref rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/r2a_launches.csv \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-state-points > gpurun_out/r2a_ncu_bench.json 2> gpurun_out/r2a_ncu.err
echo "ncu rc=$?"
