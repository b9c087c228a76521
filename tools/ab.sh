#!/bin/bash
# A/B wall-clock comparison of two builds of libb200al.so on one box: tools/ab.sh [n_images] [reps]
D=medical-image-analysis_b200/b200al
N=${1:-1024}; R=${2:-2}
for rep in $(seq $R); do
  for v in A B; do
    cp $D/lib$v.so $D/libb200al.so
    echo "== $v rep $rep"
    python tools/gpu_diag.py perf bf16x3 64 $N 1 2>&1 | grep -E "scores|conv kernels|layer  1|layer  3|layer  4|layer  8|layer 10|layer 13|layer 16|layer 18"
  done
done
