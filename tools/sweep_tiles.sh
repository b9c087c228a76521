#!/bin/bash
# Planner-knob sweep on the GPU box (per-layer us per sample of the default precision for each setting).  Builds a
# TUNING library first (-DB200AL_TUNING: the only build in which the B200AL_* environment switches are read), so run
# it on a scratch copy (gpurun) and never ship that build.  Round-2 result (profiles/r02_sweep_tiles.txt): nothing
# beyond run-to-run drift (4 % over the sweep as the GPU warms up) except the settings the planner already avoids.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
cd medical-image-analysis_b200 && B200AL_TUNING_BUILD=1 python -c "from b200al import build; build.build(force=True)" 2>&1 | tail -1; cd ..
run() { echo "== $*"; env "$@" python tools/layer_times.py --precisions mixed --out gpurun_out/sweep.jsonl 2>&1 | tail -1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        r=json.loads(l); L=r['layers']; print(r['conv_us_per_sample'], ' '.join('%s=%.2f'%(k,v['us_per_sample']) for k,v in L.items()))
    else: print(l[:200])
"; }
run X=1
run B200AL_MT2_BK=32
run B200AL_NGROUPS=1
run B200AL_NGROUPS=2
run B200AL_CW=32
run B200AL_SMEM_BUDGET=0
run B200AL_SMEM_BUDGET=116000
run X=1
