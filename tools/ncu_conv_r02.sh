#!/bin/bash
# ncu records of the conv path in the default precision (run on the GPU box):
#   gpurun_out/r02_ncu_launches.csv   launch list (gpu__time_duration) of a short bench.py run - shares of the step
#   gpurun_out/r02_ncu_full_conv.csv  --set full summary of the conv launches of one 128-sample chunk
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
B="python bench.py --pool 512 --steps 1 --warmup 1 --no-config4 --no-cpu-baseline"
$B > gpurun_out/r02_ncu_bench_plain.json 2> gpurun_out/r02_ncu_bench_plain.err || { echo "plain bench failed"; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/r02_ncu_launches.csv $B > gpurun_out/r02_ncu_launches.log 2>&1
echo "launch list: $(wc -l < gpurun_out/r02_ncu_launches.csv) lines"
python tools/ncu_targets.py conv > gpurun_out/r02_ncu_conv_plain.log 2>&1 || { echo "plain conv failed"; exit 1; }
# the second pass of the script (launches 24..46 of the conv kernels) is the warm one
timeout 1500 ncu --set full --clock-control none --import-source on -k "regex:conv_tc_kernel|conv_fold_kernel|conv_first_kernel" -s 24 -c 24 -f -o gpurun_out/r02_ncu_conv python tools/ncu_targets.py conv > gpurun_out/r02_ncu_conv.log 2>&1
python tools/ncu_summary.py gpurun_out/r02_ncu_conv.ncu-rep > gpurun_out/r02_ncu_full_conv.csv 2>> gpurun_out/r02_ncu_conv.log
ls -la gpurun_out/r02_ncu_conv.ncu-rep | awk '{print "conv report bytes:", $5}'
rm -f gpurun_out/r02_ncu_conv.ncu-rep
echo "conv summary: $(wc -l < gpurun_out/r02_ncu_full_conv.csv) rows"
