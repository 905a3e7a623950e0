#!/bin/bash
# A/B timing of library variants on conv_out / conv_in (run on the GPU box): tools/ab_co.sh VARIANT...   ("main" = the in-tree build)
for v in "$@"; do
  if [ "$v" = main ]; then unset DDIMB_LIB_PATH; else export DDIMB_LIB_PATH=$PWD/ddim_audio_b200/lib/variants/$v.so; fi
  echo "== $v"
  python tools/run_op.py conv_out1 --T 8192 --B 16 --iters 8 2>&1 | tail -1
  python tools/run_op.py conv_out --T 8192 --B 16 --iters 8 2>&1 | tail -1
  python tools/run_op.py conv_in --T 8192 --B 16 --iters 8 2>&1 | tail -1
done
