#!/bin/bash
# A/B timing of library variants on the window-kernel shapes (run on the GPU box): tools/ab_win.sh VARIANT...   ("main" = the in-tree build)
B=${B:-32}
for v in "$@"; do
  if [ "$v" = main ]; then unset DDIMB_LIB_PATH; else export DDIMB_LIB_PATH=$PWD/ddim_audio_b200/lib/variants/$v.so; fi
  echo "== $v"
  for m in 2 1; do GN_MODE=$m python tools/run_op.py conv3x3_gn --C 96 --W 64 --T 2048 --B $B --iters 8 2>&1 | tail -1; done
  python tools/run_op.py conv3x3 --C 96 --W 64 --T 2048 --B $B --iters 8 2>&1 | tail -1
  GN_MODE=2 python tools/run_op.py conv3x3_gn --C 128 --W 32 --T 1024 --B $B --iters 8 2>&1 | tail -1
  python tools/run_op.py conv3x3 --C 128 --W 32 --T 1024 --B $B --iters 8 2>&1 | tail -1
done
