"""Generate tests/golden/badge_1k.npz: the UNMODIFIED reference BADGESelector on a 1 000-image pool (container only;
needs /root/reference; about five minutes of CPU).  Usage:  python -m oracle.make_golden_badge_1k

Same construction as make_golden_badge.py (al_train's losses, batch_size=1, multiple_loss="add") at forty times the
pool: K=3 golden model, 16 labeled + 1 000 pool images of the seeded synthetic set, budget 32.  Stored: the reference
gradient embeddings [1000, 96] fp32, the names picked IN ORDER with numpy's global RNG seeded by `seed` (the
best-conditioned k-means++ run among seeds 777..976, for the reasons given in make_golden_badge.py), and that
conditioning (`gap`): an implementation whose embeddings differ by more than that need not pick the same names."""
from __future__ import annotations

import importlib
import os
import sys
import time

import numpy as np
import torch

from .kmean_ref import kmeans_plusplus_indices
from .ref_import import REFERENCE_SRC, build_reference_unet, load_reference
from .synth import make_active, synth_pool_u8
from .unet_ref import golden_state_dict

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
K, N_LAB, N_POOL, BUDGET, NP_SEED = 3, 16, 1000, 32, 777


def main():
    ref = load_reference()
    sys.path.insert(0, REFERENCE_SRC)
    try:
        al = importlib.import_module("activelearning")
        dl = importlib.import_module("losses.dice_loss")
    finally:
        sys.path.remove(REFERENCE_SRC)
    g = np.load(os.path.join(OUT, "model.npz"))
    torch.set_num_threads(os.cpu_count())
    model = build_reference_unet(K)
    model.load_state_dict(golden_state_dict(K, g))
    dice = dl.DiceLoss(num_classes=K - 1, smooth=1e-5, do_bg=True, softmax=True, batch=False, squared=False)
    sel = al.BADGESelector(dice_loss=dice, ce_loss=torch.nn.CrossEntropyLoss(), batch_size=1, num_workers=0,
                           pin_memory=False, multiple_loss="add")
    images = synth_pool_u8(N_LAB + N_POOL)
    t0 = time.time()
    names, emb = sel.cal_scores(make_active(images, N_LAB, reference=ref), model, torch.device("cpu"))
    print(f"reference cal_scores over {N_POOL} images: {time.time() - t0:.0f}s", flush=True)
    # 160 uniform draws over 1 000 cumulative weights cannot all stay 4e-3 away from a boundary (the small cases'
    # criterion): take the best-conditioned of 200 seeds and record how well conditioned that is
    seed, gap, idx = max(((sd, *kmeans_plusplus_indices(emb, BUDGET, None, np.random.RandomState(sd), return_gap=True)[::-1])
                          for sd in range(NP_SEED, NP_SEED + 200)), key=lambda t: t[1])
    np.random.seed(seed)
    picked = sel.select_next_batch(make_active(images, N_LAB, reference=ref), BUDGET, model, torch.device("cpu"))
    assert [str(names[i]) for i in idx] == list(map(str, picked))   # the k-means++ restatement agrees
    path = os.path.join(OUT, "badge_1k.npz")
    np.savez_compressed(path, emb=emb.astype(np.float32), selected=np.array(list(map(str, picked))),
                        gap=np.float64(gap), seed=np.int64(seed), n_labeled=np.int64(N_LAB), n_pool=np.int64(N_POOL))
    print("seed", seed, "gap", gap, "wrote", os.path.getsize(path), "bytes", list(picked)[:6])


if __name__ == "__main__":
    main()
