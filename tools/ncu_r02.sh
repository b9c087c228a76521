#!/bin/bash
# ncu --set full captures of the non-conv kernel families (run on the GPU box; writes gpurun_out/r02_ncu_*.ncu-rep and
# profiles-ready CSV summaries gpurun_out/r02_ncu_full_*.csv).  Usage: bash tools/ncu_r02.sh
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python tools/ncu_targets.py all > gpurun_out/r02_ncu_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/r02_ncu_plain.log; exit 1; }
run() {  # name, which, kernel regex, count
  timeout 900 ncu --set full --clock-control none --import-source on -k "regex:$3" -c "$4" -f -o "gpurun_out/r02_ncu_$1" \
      python tools/ncu_targets.py "$2" > "gpurun_out/r02_ncu_$1.log" 2>&1
  python tools/ncu_summary.py "gpurun_out/r02_ncu_$1.ncu-rep" > "gpurun_out/r02_ncu_full_$1.csv" 2>> "gpurun_out/r02_ncu_$1.log"
  rm -f "gpurun_out/r02_ncu_$1.ncu-rep"   # gpurun copies at most 64 MiB back: keep the condensed CSV only
  echo "$1: $(wc -l < gpurun_out/r02_ncu_full_$1.csv) rows"
}
run uncertainty_128 unc 'uncertainty_logits_kernel|dice_logits_kernel' 4
run kcenter kcenter 'kcenter_init_kernel|kcenter_step_kernel|kcenter_pick_kernel|normalize_rows' 12
run topk topk 'topk_' 40
run kmpp kmpp 'kmpp_' 24
run first first 'conv_first_kernel|avgpool_kernel|mc_mask' 4
