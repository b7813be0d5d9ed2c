set -x
python -m pytest tests -m gpu -q > gpurun_out/r1d_pytest.log 2>&1; tail -3 gpurun_out/r1d_pytest.log
python bench.py > gpurun_out/r1d_bench_n1.json 2> gpurun_out/r1d_bench_n1.err
python bench.py --impl reference > gpurun_out/r1d_bench_reference_arm.json 2> gpurun_out/r1d_ref.err
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r1d_smoke.log 2>&1; tail -2 gpurun_out/r1d_smoke.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r1d_launches_nt512.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extra-configs --nt 512 > gpurun_out/r1d_ncu.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:theta_kernel --launch-skip 30 --launch-count 3 -f -o gpurun_out/r1d_theta python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extra-configs > gpurun_out/r1d_ncu2.log 2>&1
ls -la gpurun_out | tail -8
