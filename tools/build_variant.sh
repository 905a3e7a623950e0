#!/bin/bash
# Build a kernel-variant copy of the library for A/B timing on the GPU box:
#   tools/build_variant.sh NAME -DDDIMB_WIDE_EPI=4 ...   ->  ddim_audio_b200/lib/variants/NAME.so  (use with DDIMB_LIB_PATH=...)
set -e
name=$1; shift
here=$(cd "$(dirname "$0")/.." && pwd)
src=$here/ddim_audio_b200/csrc; out=$here/ddim_audio_b200/lib/variants; tmp=$(mktemp -d)
mkdir -p "$out"
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo --use_fast_math -Xcompiler -fPIC -Xcompiler -O2 -cudart static"
for f in runtime gemm_host elementwise train_host train_elementwise conv_boundary denoise2d ema optim pfft2wav; do
  nvcc $FLAGS "$@" -c "$src/$f.cu" -o "$tmp/$f.o" &
done
wait
nvcc -shared -cudart static -o "$out/$name.so" "$tmp"/*.o
rm -rf "$tmp"
echo "$out/$name.so"
