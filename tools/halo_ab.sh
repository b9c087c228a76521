cd medical-image-analysis_b200 && B200AL_TUNING_BUILD=1 python -c "from b200al import build; build.build(force=True)" 2>&1 | tail -1; cd ..
B200AL_VERBOSE=1 timeout 300 python -m pytest tests/test_gpu_mixed.py -m gpu -q -x 2>&1 | grep "halo=1\|passed\|failed\|Error\|error" | sort | uniq | head -20
for h in 1 0 1 0; do echo HALO=$h; B200AL_HALO=$h timeout 200 python tools/layer_times.py --precisions mixed --out gpurun_out/lt_halo_$h.jsonl 2>&1 | tail -1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        r=json.loads(l); L=r['layers']; print(r['conv_us_per_sample'], r['max_abs_logit_err_vs_reference'], ' '.join('%s=%.2f'%(k,L[k]['us_per_sample']) for k in ['enc2.1','enc3.1','enc4.1','dec0.0','dec0.1','dec1.0','dec1.1']))
    else: print(l[:300])
"; done
