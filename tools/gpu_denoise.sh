#!/bin/bash
# GPU pass for the denoise_2d post-filter: its parity tests, a CUDA-event timing next to torch-eager, an ncu launch list
# and one --set full capture of its kernels (summarised on the box).
set -u
R=${ROUND:-r01}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_denoise.py -m gpu -x -q > gpurun_out/pytest_denoise.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/pytest_denoise.log
timeout 600 python tools/bench_denoise.py > gpurun_out/denoise_bench.json 2> gpurun_out/denoise_bench.err; echo "bench rc=$?"
cat gpurun_out/denoise_bench.json; tail -3 gpurun_out/denoise_bench.err
SMALL="python tools/bench_denoise.py --B 16 --T 8192 --no-torch --iters 2"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/denoise_launches.csv $SMALL > gpurun_out/ncu_dn_list.log 2>&1
echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none -k regex:'row_median|col_hist|col_scan|denoise_iter|denoise_final|col_finalize' -s 28 -c 14 -f -o gpurun_out/prof_${R}_denoise $SMALL > gpurun_out/ncu_dn_full.log 2>&1
echo "ncu full rc=$?"
python tools/ncu_summarize.py launches gpurun_out/denoise_launches.csv gpurun_out/${R}_denoise_launches_summary.csv "$SMALL (-c 400)"
python tools/ncu_summarize.py full gpurun_out/prof_${R}_denoise.ncu-rep gpurun_out/${R}_denoise_ncu_full_summary.csv
rm -f gpurun_out/prof_${R}_denoise.ncu-rep
cat gpurun_out/${R}_denoise_launches_summary.csv
