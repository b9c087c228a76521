"""One representative launch of every non-conv kernel family at the sizes the BASELINE configs use, for
`ncu --set full`.  Run plain first, then under ncu with a kernel-name filter:

  python tools/ncu_targets.py                       # sanity: exits 0, prints the sizes
  ncu --set full --clock-control none --import-source on -k regex:'uncertainty_logits|dice_logits' -c 4 \\
      -o gpurun_out/r02_unc python tools/ncu_targets.py unc
  ... (tools/ncu_r02.sh runs the whole list and condenses each report with tools/ncu_summary.py)

Sizes: uncertainty reduction = 128 images x T=8 x K=3 fp32 logits (805 MB, the launch bench.py times);
k-center = 100 000 pool rows + 500 labeled, D=512 (config 3: 205.7 MB per step); top-k = 100 000 scores, k=5000
(the hybrid pre-filter); k-means++ = 100 000 x 512, 6 centres; first conv / embedding tail = 64 images.
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "medical-image-analysis_b200")]


def main(which):
    import numpy as np
    import b200al
    from bench_fixtures import StateDictModel, seeded_state_dict, synthetic_pool_u8
    dev = torch.device("cuda:0")
    eng = b200al.Engine(dev, 3, max_samples=64)
    g = torch.Generator(device="cuda").manual_seed(0)
    if which in ("all", "unc"):
        lg = torch.randn((128, 8, 3, 256, 256), device=dev, generator=g)
        for mode in ("entropy", "dice"):
            for _ in range(2):
                eng.uncertainty_from_logits(lg, mode, smooth=1e-8 if mode == "entropy" else 1e-5)
        del lg
    if which in ("all", "kcenter"):
        feats = torch.randn((100_500, 512), device=dev, generator=g) * 0.3 + 0.1
        eng.kcenter(feats, 500, 6, "l2")
        eng.kcenter(feats, 500, 6, "cosine")
        del feats
    if which in ("all", "topk"):
        sc = torch.randn(100_000, device=dev, generator=g)
        eng.topk(sc, 5000)
        eng.topk(sc, 500)
    if which in ("all", "kmpp"):
        X = torch.randn((100_000, 512), device=dev, generator=g)
        rs = np.random.RandomState(0)
        k, trials = 6, 2 + int(np.log(6))
        w = np.ones(100_000, np.float32)
        uni = np.stack([rs.uniform(size=trials) for _ in range(k - 1)])
        eng.kmeans_pp(X, torch.from_numpy(w), 123, torch.from_numpy(uni), k, trials)
        del X
    if which == "conv":   # the 23 conv launches of ONE 128-sample chunk in the default (mixed) precision
        e2 = b200al.Engine(dev, 3, max_samples=128)
        e2.load_model(StateDictModel(seeded_state_dict(3)))
        imgs = synthetic_pool_u8(128).to(dev)
        for _ in range(2):
            e2.scores(imgs, "entropy")
    if which in ("all", "first"):
        eng.load_model(StateDictModel(seeded_state_dict(3)))
        imgs = synthetic_pool_u8(64).to(dev)
        eng.scores(imgs, "entropy")
        eng.embed(imgs)
    torch.cuda.synchronize()
    print("ncu_targets", which, "done")


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else "all")
