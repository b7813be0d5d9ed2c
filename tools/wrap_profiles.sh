# Round-end evidence run (one GPU): full GPU test suite, both bench arms, smoke, ncu launch list and --set full captures.
# Usage on the GPU box:  bash tools/wrap_profiles.sh <tag>     (outputs under gpurun_out/<tag>_*)
tag=${1:-r1e}
set -x
python -m pytest tests -m gpu -q > gpurun_out/${tag}_pytest.log 2>&1; tail -3 gpurun_out/${tag}_pytest.log
python bench.py > gpurun_out/${tag}_bench_n1.json 2> gpurun_out/${tag}_bench_n1.err
python bench.py --impl reference > gpurun_out/${tag}_bench_reference_arm.json 2> gpurun_out/${tag}_ref.err
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/${tag}_smoke.log 2>&1; tail -2 gpurun_out/${tag}_smoke.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/${tag}_launches_nt512.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extra-configs --nt 512 > gpurun_out/${tag}_ncu.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:theta_kernel --launch-skip 25 --launch-count 3 -f -o gpurun_out/${tag}_theta python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extra-configs > gpurun_out/${tag}_ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:theta_setup --launch-skip 25 --launch-count 3 -f -o gpurun_out/${tag}_setup python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extra-configs > gpurun_out/${tag}_ncu4.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:mesh_level --launch-skip 25 --launch-count 1 -f -o gpurun_out/${tag}_mesh python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extra-configs > gpurun_out/${tag}_ncu3.log 2>&1
ls -la gpurun_out | grep ${tag}
