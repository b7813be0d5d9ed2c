#!/bin/bash
mkdir -p gpurun_out
for E in 100 1000; do
  ncu --set full --import-source on --clock-control none --kernel-name regex:q_xsec --launch-count 1 -f -o gpurun_out/qp_$E python tools/qprof.py $E > /dev/null 2> gpurun_out/qp_$E.err
  { python tools/ncu_digest.py gpurun_out/qp_$E.ncu-rep 40; echo; python tools/ncu_lines.py gpurun_out/qp_$E.ncu-rep 45; } > gpurun_out/qp_${E}_digest.txt 2>&1
  rm -f gpurun_out/qp_$E.ncu-rep
done
