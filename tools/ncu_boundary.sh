set -e
python -m pytest tests/test_gpu_train_kernels.py -x -q -k "conv_in_out_wgrad" 2>&1 | tail -3
python bench.py --mode train --steps 5 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
t=d.get('train', d)
print('train', t.get('samples_per_s'), t.get('ms_per_step'))
for k in t['kernels'][:40]:
    if 'wgrad' in k['name'] and ('conv_in' in k['name'] or 'conv_out' in k['name']): print(k)
"
python tools/run_op.py conv_out1 --T 8192 --B 16 --iters 3 | tail -1
python tools/run_op.py conv_in --T 8192 --B 16 --iters 3 | tail -1
ncu --set full --clock-control none --import-source on -k regex:conv_out_tc -c 1 -s 1 -o gpurun_out/r02_convout -f python tools/run_op.py conv_out1 --T 8192 --B 16 --iters 2 > gpurun_out/ncu_co.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:conv3x3_wide -c 1 -s 1 -o gpurun_out/r02_convin -f python tools/run_op.py conv_in --T 8192 --B 16 --iters 2 > gpurun_out/ncu_ci.log 2>&1
