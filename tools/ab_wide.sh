#!/bin/bash
# A/B timing of library variants on the wide-conv shapes (run on the GPU box): tools/ab_wide.sh VARIANT...   ("main" = the in-tree build)
for v in "$@"; do
  if [ "$v" = main ]; then unset DDIMB_LIB_PATH; else export DDIMB_LIB_PATH=$PWD/ddim_audio_b200/lib/variants/$v.so; fi
  echo "== $v"
  for m in 2 1; do GN_MODE=$m python tools/run_op.py conv3x3_gn --C 32 --W 256 --T 8192 --B 16 --iters 8 2>&1 | tail -1; done
  python tools/run_op.py conv3x3 --C 32 --W 256 --T 8192 --B 16 --iters 8 2>&1 | tail -1
  python tools/run_op.py conv3x3_gn --C 64 --W 128 --T 4096 --B 32 --iters 8 2>&1 | tail -1
done
