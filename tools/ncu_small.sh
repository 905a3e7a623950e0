set -e
# ncu captures of the level-0 / level-2 / level-5 convolutions at the small sampling batch x[4,2,1024,256]
GN_MODE=2 python tools/run_op.py conv3x3_gn --C 32 --W 256 --T 1024 --B 4 --iters 3 | tail -1
ncu --set full --clock-control none --import-source on -k regex:conv3x3_wide -c 1 -s 2 -o gpurun_out/r02_small_wide32 -f python tools/run_op.py conv3x3_gn --C 32 --W 256 --T 1024 --B 4 --iters 2 > gpurun_out/ncu_s1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:conv3x3_window -c 1 -s 2 -o gpurun_out/r02_small_win96 -f python tools/run_op.py conv3x3_gn --C 96 --W 64 --T 256 --B 4 --iters 2 > gpurun_out/ncu_s2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:umma_gemm -c 1 -s 2 -o gpurun_out/r02_small_c256 -f python tools/run_op.py conv3x3 --C 256 --W 8 --T 32 --B 4 --iters 2 > gpurun_out/ncu_s3.log 2>&1
