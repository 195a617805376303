#!/bin/bash
# One `ncu --set full` capture per HBM-bound kernel of scripts/bench_hbm.py (fourth launch of each), summarised with
# scripts/ncu_summary.py into gpurun_out/r02_hbm_kernels_summary.txt.  Run on the GPU box: bash scripts/ncu_hbm.sh
set -u
out=gpurun_out/r02_hbm_kernels_summary.txt
: > $out
for k in unary binary maxpool_nhwc4 softmax_rows layer_norm_rows depthwise3x3_tma; do
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:$k -s 3 -c 1 -f -o gpurun_out/r02_hbm_$k \
    python scripts/bench_hbm.py > gpurun_out/r02_hbm_$k.log 2>&1
  echo "==== $k" >> $out
  python scripts/ncu_summary.py gpurun_out/r02_hbm_$k.ncu-rep 12 >> $out 2>&1
done
rm -f gpurun_out/r02_hbm_*.ncu-rep
