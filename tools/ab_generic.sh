#!/bin/bash
# timings of the generic / window / up / down convolution kernels at the bench shapes (B = 64, T = 8192 at level 0)
B=${B:-64}
python tools/run_op.py conv3x3 --C 96 --W 64 --T 2048 --B $B --iters 5 | tail -1
python tools/run_op.py conv3x3 --C 128 --W 32 --T 1024 --B $B --iters 5 | tail -1
python tools/run_op.py conv3x3 --C 192 --W 16 --T 512 --B $B --iters 5 | tail -1
python tools/run_op.py conv3x3 --C 256 --W 8 --T 256 --B $B --iters 5 | tail -1
python tools/run_op.py down --C 32 --Cout 64 --W 256 --T 8192 --B $B --iters 3 | tail -1
python tools/run_op.py down --C 64 --Cout 96 --W 128 --T 4096 --B $B --iters 3 | tail -1
python tools/run_op.py down --C 96 --Cout 128 --W 64 --T 2048 --B $B --iters 3 | tail -1
python tools/run_op.py down --C 128 --Cout 192 --W 32 --T 1024 --B $B --iters 3 | tail -1
python tools/run_op.py up --C 64 --Cout 32 --W 128 --T 4096 --B $B --iters 3 | tail -1
python tools/run_op.py up --C 96 --Cout 64 --W 64 --T 2048 --B $B --iters 3 | tail -1
python tools/run_op.py up --C 128 --Cout 96 --W 32 --T 1024 --B $B --iters 3 | tail -1
python tools/run_op.py up --C 192 --Cout 128 --W 16 --T 512 --B $B --iters 3 | tail -1
python tools/run_op.py up --C 256 --Cout 192 --W 8 --T 256 --B $B --iters 3 | tail -1
python tools/run_op.py conv_in --T 8192 --B $B --iters 3 | tail -1
python tools/run_op.py conv_out --T 8192 --B $B --iters 3 | tail -1
