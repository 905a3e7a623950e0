#!/bin/bash
# One GPU-box pass: parity tests, the default bench line, ncu launch list, --set full captures of the top kernels (small shape)
# and one full-shape capture of the dominant kernel (DRAM traffic per launch for bench.py's roofline.traffic).
set -u
R=${ROUND:-r01}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
python bench.py --mode train > gpurun_out/bench_train.json 2> gpurun_out/bench_train.err; echo "train bench rc=$?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "reference arm rc=$?"
SMALL="python bench.py --steps 2 --warmup 1 --batch 8 --t-size 2048 --no-cpu-baseline --no-e2e"
$SMALL > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/launches.csv $SMALL > gpurun_out/ncu_list.log 2>&1
echo "ncu list rc=$?"
$SMALL > gpurun_out/plain2.log 2>&1 && \
ncu --set full --clock-control none -k regex:'gn_apply|conv3x3_wide_kernel|conv3x3_window_kernel|umma_gemm_kernel|conv_up_wide|conv_down_wide|conv_out_tc|conv_in_strip' -s 150 -c 48 -f -o gpurun_out/prof_${R} $SMALL > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
FULL="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
ncu --set full --clock-control none -k regex:conv3x3_wide_kernel -s 4 -c 2 -f -o gpurun_out/prof_${R}_fullshape $FULL > gpurun_out/ncu_fullshape.log 2>&1
echo "ncu full-shape rc=$?"
# summarise on the box (the reports are too large to travel back: gpurun_out is capped at 64 MiB)
python tools/ncu_summarize.py launches gpurun_out/launches.csv gpurun_out/${R}_launches_summary.csv "$SMALL (-c 1200)"
python tools/ncu_summarize.py full gpurun_out/prof_${R}.ncu-rep gpurun_out/${R}_ncu_full_summary.csv
python tools/ncu_summarize.py full gpurun_out/prof_${R}_fullshape.ncu-rep gpurun_out/${R}_ncu_fullshape_summary.csv
rm -f gpurun_out/prof_${R}.ncu-rep gpurun_out/launches.csv
ls -la gpurun_out
