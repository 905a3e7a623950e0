#!/bin/bash
# Multi-GPU pass on ONE box: the bench line at N ranks (batch-sharded sampling, no collective) and the data-parallel train line
# (NCCL all-reduce of the flat gradient buffer).  N from $NGPU (default 8).
set -u
N=${NGPU:-8}
R=${ROUND:-r01}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 420 $TR --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/${R}_bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "sample N=$N rc=$?"
cut -c1-700 gpurun_out/${R}_bench_n$N.json; tail -3 gpurun_out/bench_n$N.err
timeout 420 $TR --master-port 29512 bench.py --gpus $N --mode train --steps 5 --warmup 3 > gpurun_out/${R}_bench_train_n$N.json 2> gpurun_out/bench_train_n$N.err; echo "train N=$N rc=$?"
cut -c1-700 gpurun_out/${R}_bench_train_n$N.json; tail -3 gpurun_out/bench_train_n$N.err
