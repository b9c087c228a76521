"""Generate tests/golden/badge.npz by running the UNMODIFIED reference BADGESelector here (container only; needs
/root/reference).  Usage:  python -m oracle.make_golden_badge

Golden model (tests/golden/model.npz) in the reference UNet, losses exactly as al_train builds them
(al_trainer.py:782-798: DiceLoss(num_classes, smooth=1e-5, do_bg=True, softmax=True, batch=False, squared=False) and
torch.nn.CrossEntropyLoss()), batch_size=1, multiple_loss="add" (al_trainer.py:871-879).  Stored per case:
  emb/<case>       reference gradient embeddings [n_pool, K*32] fp32 (`cal_scores`)
  selected/<case>  names picked by `select_next_batch` IN PICK ORDER, numpy's global RNG seeded with seed/<case> before
  gap/<case>       conditioning of that k-means++ run (oracle.kmean_ref, return_gap).  The seed is the first one from
                   NP_SEED upwards whose run is well conditioned (gap > 4e-3).  Two reasons: with a near-tie between
                   two candidates' potentials sklearn's own pick depends on its fp32 BLAS summation order (seed 777 on K3:
                   two candidates tie to 1e-8 relative); and pseudo-labels are an argmax, so logits that differ within the
                   north star's tolerance flip the label of a few near-tie pixels, which moves a gradient embedding by up
                   to ~1e-3 of its scale - any implementation that is not bit-identical in the logits needs that margin.
plus emb/K3/nobg: embeddings with DiceLoss(do_bg=False) on 6 images (not reachable from al_train, supported by the ABI).
"""
from __future__ import annotations

import importlib
import os
import sys

import numpy as np
import torch

from .kmean_ref import kmeans_plusplus_indices
from .ref_import import REFERENCE_SRC, build_reference_unet, load_reference
from .synth import make_active, synth_pool_u8
from .unet_ref import golden_state_dict

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
NP_SEED = 777
# (key, K, n_labeled, n_pool, budget, do_bg)
CASES = [("K3", 3, 8, 24, 6, True), ("K2", 2, 8, 24, 6, True), ("K3/nobg", 3, 8, 6, 0, False)]


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
    out = {}
    for key, K, n_lab, n_pool, budget, do_bg in CASES:
        model = build_reference_unet(K)
        model.load_state_dict(golden_state_dict(K, g))
        dice = dl.DiceLoss(num_classes=K - 1, smooth=1e-5, do_bg=do_bg, softmax=True, batch=False, squared=False)
        sel = al.BADGESelector(dice_loss=dice, ce_loss=torch.nn.CrossEntropyLoss(), batch_size=1, num_workers=0,
                               pin_memory=False, multiple_loss="add")
        images = synth_pool_u8(n_lab + n_pool)
        names, emb = sel.cal_scores(make_active(images, n_lab, reference=ref), model, torch.device("cpu"))
        out["emb/" + key] = emb.astype(np.float32)
        if budget:
            seed = NP_SEED
            while True:
                idx, gap = kmeans_plusplus_indices(emb, budget, None, np.random.RandomState(seed), return_gap=True)
                if gap > 4e-3:
                    break
                seed += 1
            np.random.seed(seed)
            picked = sel.select_next_batch(make_active(images, n_lab, reference=ref), budget, model, torch.device("cpu"))
            out["selected/" + key] = np.array(list(map(str, picked)))
            assert [str(names[i]) for i in idx] == list(map(str, picked))  # the k-means++ restatement agrees
            out["gap/" + key] = np.float64(gap)
            out["seed/" + key] = np.int64(seed)
            print(key, emb.shape, list(picked), f"seed {seed} gap {gap:.2e}")
    np.savez_compressed(os.path.join(OUT, "badge.npz"), **out)
    print("wrote", os.path.getsize(os.path.join(OUT, "badge.npz")), "bytes")


if __name__ == "__main__":
    main()
