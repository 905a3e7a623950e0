set -e
python tools/run_op.py conv3x3_gn --C 96 --W 64 --T 2048 --B 16 --iters 3 | tail -1
python tools/run_op.py up --C 96 --Cout 64 --W 64 --T 2048 --B 16 --iters 3 | tail -1
python tools/run_op.py down --C 64 --Cout 96 --W 128 --T 4096 --B 16 --iters 3 | tail -1
ncu --set full --clock-control none --import-source on -k regex:conv3x3_window -c 1 -s 2 -o gpurun_out/r02_win96 -f python tools/run_op.py conv3x3_gn --C 96 --W 64 --T 2048 --B 16 --iters 2 > gpurun_out/ncu_win.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:umma_gemm -c 2 -s 2 -o gpurun_out/r02_up96 -f python tools/run_op.py up --C 96 --Cout 64 --W 64 --T 2048 --B 16 --iters 2 > gpurun_out/ncu_up.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:umma_gemm -c 1 -s 1 -o gpurun_out/r02_down64 -f python tools/run_op.py down --C 64 --Cout 96 --W 128 --T 4096 --B 16 --iters 2 > gpurun_out/ncu_down.log 2>&1
ls -la gpurun_out/*.ncu-rep
