#!/bin/bash
# A/B of the gn_apply variants at the bench shapes (DDIMB_GN_OLD=1: 2-D grid kernel, 2: balanced ranges, unset: C >= 96 -> ranges).
for shape in "96 64 2048" "128 32 1024" "192 16 512" "64 128 4096"; do
  set -- $shape
  for op in ${OPS:-gn}; do
    for cfg in "DDIMB_GN_OLD=1" "DDIMB_GN_OLD=2 DDIMB_GN_UNR=4" "DDIMB_GN_OLD=2 DDIMB_GN_UNR=8" "DDIMB_GN_OLD=1 DDIMB_GN_UNR=8"; do
      echo -n "C=$1 W=$2 T=$3 $op [$cfg]: "
      env $cfg python tools/run_op.py $op --C $1 --W $2 --T $3 --B 64 --iters 5 2>&1 | tail -1
    done
  done
done
