#!/bin/bash
# A/B of the row-halo pipeline stages on the GPU box, in a TUNING build (never ship that build):
#   B200AL_HALO  (default 1)  fp16 stride-1 3x3 layers;  B200AL_HALO3 (default: see b200al.cu)  bf16x3 N = 64 super-tile
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
(cd medical-image-analysis_b200 && B200AL_TUNING_BUILD=1 python -c "from b200al import build; build.build(force=True)" 2>&1 | tail -1)
B200AL_HALO3=1 timeout 300 python -m pytest tests/test_gpu_layers.py tests/test_gpu_mixed.py -m gpu -q -x 2>&1 | tail -3
for h in 1 0 1 0; do echo HALO3=$h; B200AL_HALO3=$h timeout 200 python tools/layer_times.py --precisions mixed --out gpurun_out/lt_halo3_$h.jsonl 2>&1 | tail -1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        r=json.loads(l); L=r['layers']; print(r['conv_us_per_sample'], r['max_abs_logit_err_vs_reference'], ' '.join('%s=%.2f'%(k,L[k]['us_per_sample']) for k in ['dec2.0','dec2.1','dec1.0','enc2.1']))
    else: print(l[:300])
"; done
