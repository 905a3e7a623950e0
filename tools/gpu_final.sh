#!/bin/bash
# Final pass of a round on one GPU box: parity tests, default bench line, ncu launch list at the full shape, one --set full capture of the
# dominant kernel (each ncu command only after the same command has exited 0 without ncu).
set -u
R=${ROUND:-r02}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "reference arm rc=$?"
FULL="python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-train --no-small-batch --no-parity"
$FULL > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/launches.csv $FULL > gpurun_out/ncu_list.log 2>&1
echo "ncu list rc=$?"
$FULL > gpurun_out/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:conv3x3_wide_kernel -s 0 -c 2 -f -o gpurun_out/prof_${R}_fullshape $FULL > gpurun_out/ncu_fullshape.log 2>&1
echo "ncu full-shape rc=$?"
python tools/ncu_summarize.py launches gpurun_out/launches.csv gpurun_out/${R}_launches_fullshape_summary.csv "$FULL (x[64,2,8192,256])"
python tools/ncu_summarize.py full gpurun_out/prof_${R}_fullshape.ncu-rep gpurun_out/${R}_ncu_fullshape_c32_summary.csv
rm -f gpurun_out/launches.csv
ls -la gpurun_out | head -30
