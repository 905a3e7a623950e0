#!/bin/bash
# Bottleneck isolation of the wide conv kernel via DDIMB_WIDE_DBG bits (2: no epilogue, 4: no MMA, 8: L2-resident loads, 16: no stores, 32: no store staging)
for d in ${DBGS:-0 2 4 6 16}; do
  echo "== dbg $d"
  DDIMB_WIDE_DBG=$d python tools/run_op.py conv3x3_gn --C 32 --W 256 --T 8192 --B 8 --iters 5 2>&1 | tail -1
  DDIMB_WIDE_DBG=$d python tools/run_op.py conv3x3_gn --C 64 --W 128 --T 4096 --B 16 --iters 5 2>&1 | tail -1
done
