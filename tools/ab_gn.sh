#!/bin/bash
# A/B timing of library variants on gn_apply mode 2 (GN2 + residual) at the bench shapes: tools/ab_gn.sh VARIANT...   ("main" = the in-tree build)
for v in "$@"; do
  if [ "$v" = main ]; then unset DDIMB_LIB_PATH; else export DDIMB_LIB_PATH=$PWD/ddim_audio_b200/lib/variants/$v.so; fi
  echo "== $v"
  for shape in "96 64 2048" "128 32 1024" "192 16 512" "256 8 256"; do
    set -- $shape
    python tools/run_op.py gn2 --C $1 --W $2 --T $3 --B 64 --iters 8 2>&1 | tail -1
  done
done
