#!/bin/bash
# Bottleneck isolation of the wide conv kernel via DDIMB_WIDE_DBG bits (2: no epilogue, 4: no MMA, 8: L2-resident loads, 16: no stores)
for d in 0 2 4 6 14; do
  echo "== dbg $d"
  DDIMB_WIDE_DBG=$d python tools/run_op.py conv3x3 --C 32 --W 256 --T 8192 --B 8 --iters 5 2>&1 | tail -1
  DDIMB_WIDE_DBG=$d python tools/run_op.py conv3x3 --C 64 --W 128 --T 4096 --B 16 --iters 5 2>&1 | tail -1
done
for d in 0 2; do
DDIMB_WIDE_DBG=$d ncu --set full --clock-control none --import-source on -k regex:conv3x3_wide -s 2 -c 1 -o gpurun_out/wide32_dbg$d -f python tools/run_op.py conv3x3 --C 32 --W 256 --T 8192 --B 8 --iters 3 > /dev/null 2>&1
done
ls gpurun_out
