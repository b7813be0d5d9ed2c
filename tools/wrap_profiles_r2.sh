#!/bin/bash
# Round-2 evidence run (one GPU): GPU test suite, both bench arms, smoke, ncu launch list of the headline step, FP64 op counters,
# --set full captures of the hot kernels. Usage on the GPU box: bash tools/wrap_profiles_r2.sh <tag>  (outputs gpurun_out/<tag>_*)
tag=${1:-r2}
mkdir -p gpurun_out
B="--no-cpu-baseline --no-state-points --no-extra-configs"
python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/${tag}_pytest_gpu.log 2>&1; tail -2 gpurun_out/${tag}_pytest_gpu.log
python bench.py --impl reference > gpurun_out/${tag}_bench_reference_arm.json 2> gpurun_out/${tag}_ref.err; echo "ref arm rc=$?"
python bench.py > gpurun_out/${tag}_bench_n1.json 2> gpurun_out/${tag}_bench_n1.err; echo "bench rc=$?"
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/${tag}_smoke.log 2>&1; tail -1 gpurun_out/${tag}_smoke.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/${tag}_launches.csv python bench.py --steps 1 --warmup 3 $B > gpurun_out/${tag}_ncu_launches.log 2>&1
M=smsp__sass_thread_inst_executed_op_dfma_pred_on.sum,smsp__sass_thread_inst_executed_op_dmul_pred_on.sum,smsp__sass_thread_inst_executed_op_dadd_pred_on.sum,gpu__time_duration.sum
ncu --metrics $M --clock-control none --kernel-name regex:"theta|mesh_level" -c 400 --csv --log-file gpurun_out/${tag}_fp64_ops.csv python bench.py --steps 1 --warmup 3 $B > gpurun_out/${tag}_fp64_ops_bench.json 2> /dev/null
for k in theta_setup theta_kernel mesh_level; do
  ncu --set full --import-source on --clock-control none --kernel-name regex:$k --launch-skip 4 --launch-count 1 -f -o gpurun_out/${tag}_$k python bench.py --steps 1 --warmup 3 $B > /dev/null 2> gpurun_out/${tag}_ncu_$k.err
done
ncu --set full --import-source on --clock-control none --kernel-name regex:mesh_level --launch-skip 3 --launch-count 1 -f -o gpurun_out/${tag}_mesh_level_p4 python bench.py --steps 1 --warmup 3 $B > /dev/null 2>> gpurun_out/${tag}_ncu_mesh_level.err
ncu --set full --import-source on --clock-control none --kernel-name regex:state_kernel --launch-skip 2 --launch-count 1 -f -o gpurun_out/${tag}_state python tools/profile_state.py > /dev/null 2> gpurun_out/${tag}_ncu_state.err
ncu --set full --import-source on --clock-control none --kernel-name regex:rdf_kernel --launch-skip 2 --launch-count 1 -f -o gpurun_out/${tag}_rdf python tools/profile_state.py > /dev/null 2>> gpurun_out/${tag}_ncu_state.err
# EoS and quantum kernels: FP64 op counters of every launch, one full capture each
ncu --metrics $M --clock-control none --kernel-name regex:"eos_" --csv --log-file gpurun_out/${tag}_eos_fp64_ops.csv python tools/profile_eos.py > /dev/null 2>&1
ncu --set full --import-source on --clock-control none --kernel-name regex:eos_hess --launch-skip 2 --launch-count 1 -f -o gpurun_out/${tag}_eos_hess python tools/profile_eos.py > /dev/null 2> gpurun_out/${tag}_ncu_eos.err
ncu --metrics $M --clock-control none --kernel-name regex:"q_xsec" --csv --log-file gpurun_out/${tag}_quantum_fp64_ops.csv python tools/profile_quantum.py > gpurun_out/${tag}_quantum_ops.log 2>&1
ncu --set full --import-source on --clock-control none --kernel-name regex:q_xsec --launch-skip 2 --launch-count 1 -f -o gpurun_out/${tag}_q_xsec python tools/profile_quantum.py > /dev/null 2> gpurun_out/${tag}_ncu_quantum.err
# digest the captures here (ncu reads them on the box) and drop the reports: gpurun brings back at most 64 MiB
for k in theta_setup theta_kernel mesh_level mesh_level_p4 state rdf eos_hess q_xsec; do
  rep=gpurun_out/${tag}_$k.ncu-rep
  [ -s $rep ] || continue
  ncu -i $rep --page raw --csv > gpurun_out/${tag}_ncu_${k}_raw.csv 2>/dev/null
  { python tools/ncu_digest.py $rep 50; echo; python tools/ncu_lines.py $rep 30; } > gpurun_out/${tag}_ncu_${k}_digest.txt 2>&1
  rm -f $rep
done
ls -la gpurun_out | grep ${tag}_
