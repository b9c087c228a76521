"""Per-precision record on one B200: logit / score error against the reference-frozen goldens and per-layer kernel
time (CUDA events around every conv launch, b200al_set_timing) of an eval-mode scoring pass.

  python tools/layer_times.py [--samples 256] [--out gpurun_out/r02_layer_times.jsonl]
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "medical-image-analysis_b200")]

NAMES = ["enc0.1", "enc1.0", "enc1.1", "enc2.0", "enc2.1", "enc3.0", "enc3.1", "enc4.0", "enc4.1", "up0", "dec0.0",
         "dec0.1", "up1", "dec1.0", "dec1.1", "up2", "dec2.0", "dec2.1", "up3", "dec3.0", "dec3.1"]
GF = [1.208, 0.604, 1.208, 0.604, 1.208, 0.604, 1.208, 0.604, 1.208, 0.268, 2.416, 1.208, 0.268, 2.416, 1.208, 0.268,
      2.416, 1.208, 0.268, 2.416, 1.208]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--samples", type=int, default=256)
    ap.add_argument("--reps", type=int, default=4)
    ap.add_argument("--precisions", default="bf16x3,mixed,mixed-deep")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "r02_layer_times.jsonl"))
    args = ap.parse_args()
    import b200al
    from bench_fixtures import seeded_state_dict, synthetic_pool_u8
    dev = torch.device("cuda:0")
    g = np.load(os.path.join(ROOT, "tests", "golden", "pool_K3.npz"))
    L, n = int(g["n_labeled"]), int(g["n_pool"])
    sd = seeded_state_dict(3)
    pool = synthetic_pool_u8(L + n)[L:].to(dev)       # == oracle.synth images (bench_fixtures restates the generator)
    big = synthetic_pool_u8(args.samples).to(dev)
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    for prec in args.precisions.split(","):
        eng = b200al.Engine(dev, 3, max_samples=args.samples, precision=prec)
        eng.load_model(sd)
        _s, logits = eng.scores(pool[:2], "entropy", return_logits=True)
        logit_err = float(np.abs(logits.cpu().numpy() - g["probe_logits"]).max())
        rel = {m: float((np.abs(eng.scores(pool, m).cpu().numpy() - g[f"scores/{m}"]) / np.abs(g[f"scores/{m}"])).max())
               for m in ("entropy", "margin", "confidence")}
        for _ in range(2):
            eng.scores(big, "entropy")
        eng.set_timing(True)
        for _ in range(args.reps):
            eng.scores(big, "entropy")
        ms, launches, flops = eng.conv_time_ms()
        per = [t * 1e3 / (args.reps * args.samples) for t in eng.layer_times_ms()]   # us per sample
        eng.set_timing(False)

        def whole(fn):   # us per sample of the whole call (first conv, conv layers, reduction), CUDA events
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            fn()
            e0.record()
            for _ in range(args.reps):
                fn()
            e1.record()
            torch.cuda.synchronize()
            return round(e0.elapsed_time(e1) * 1e3 / (args.reps * args.samples), 2)

        total_scores = whole(lambda: eng.scores(big, "entropy"))
        total_embed = whole(lambda: eng.embed(big))
        rec = {"precision": prec, "samples": args.samples, "scores_call_us_per_sample": total_scores,
               "embed_call_us_per_sample": total_embed, "max_abs_logit_err_vs_reference": logit_err,
               "max_rel_score_err_vs_reference": rel, "conv_us_per_sample": round(sum(per), 2),
               "conv_tflops_algorithmic": round(flops / (ms / 1e3) / 1e12, 1),
               "layers": {nm: {"us_per_sample": round(t, 3), "tflops": round(gf * 1e9 / (t * 1e-6) / 1e12, 1) if t > 0 else None}
                          for nm, t, gf in zip(NAMES, per, GF)}}
        print(json.dumps(rec), flush=True)
        with open(args.out, "a") as f:
            f.write(json.dumps(rec) + "\n")
        eng.close()


if __name__ == "__main__":
    main()
