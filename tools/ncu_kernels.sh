#!/bin/bash
# One ncu --set full capture each of the secondary kernels at their bench shapes (after the same command has run once without ncu),
# summarised on the box into one CSV (the reports themselves are too large to keep).
set -u
mkdir -p gpurun_out/ncu_k
cap() {  # name, kernel regex, skip, run_op args...
  local name=$1 rx=$2 skip=$3; shift 3
  python tools/run_op.py "$@" --iters 2 > /dev/null 2>&1 || { echo "$name: plain run failed"; return; }
  ncu --set full --clock-control none -k regex:$rx -s $skip -c 1 -f -o gpurun_out/ncu_k/$name python tools/run_op.py "$@" --iters 2 > gpurun_out/ncu_k/$name.log 2>&1
  python tools/ncu_summarize.py full gpurun_out/ncu_k/$name.ncu-rep gpurun_out/ncu_k/$name.csv
}
GN_MODE=2 cap win96 conv3x3_window 1 conv3x3_gn --C 96 --W 64 --T 2048 --B 16
GN_MODE=2 cap win128 conv3x3_window 1 conv3x3_gn --C 128 --W 32 --T 1024 --B 16
GN_MODE=2 cap wide64 conv3x3_wide 1 conv3x3_gn --C 64 --W 128 --T 4096 --B 16
cap convout conv_out_tc 1 conv_out1 --T 8192 --B 16
cap up96 umma_gemm 2 up --C 96 --Cout 64 --W 64 --T 2048 --B 16
cap wgrad32 wgrad_wide 1 wgrad --C 32 --W 256 --T 1024 --B 32
cap wgrad64 wgrad_wide 1 wgrad --C 64 --W 128 --T 512 --B 32
cap wgradud wgrad_updown 1 wgrad_down --C 32 --Cout 64 --W 256 --T 1024 --B 32
cap gn2 gn_apply 1 gn2 --C 32 --W 256 --T 8192 --B 16
{ head -2 gpurun_out/ncu_k/win96.csv; for f in gpurun_out/ncu_k/*.csv; do tail -n +3 $f; done; } > gpurun_out/r02_ncu_kernels_summary.csv
rm -f gpurun_out/ncu_k/*.ncu-rep
cat gpurun_out/r02_ncu_kernels_summary.csv | cut -c1-260
