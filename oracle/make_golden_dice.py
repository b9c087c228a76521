"""tests/golden/dice.npz: dice-uncertainty scores produced with the reference's own DiceLoss class (container only).

  python -m oracle.make_golden_dice

Dice-uncertainty is named by BASELINE.json's north star and defined in SURVEY 8c; the reference has no selector for
it, so the score is COMPOSED ("parity unpinned by the reference" as a selector), but every arithmetic step is the
reference's: reference UNet logits -> `DiceLoss(num_classes=K-1, smooth=1e-5, do_bg=True, softmax=True)` (the
instance al_trainer.py:785-791 builds) called per image on (logits, argmax of its softmax).  Pools = those of
pool_K2.npz / pool_K3.npz.
"""
from __future__ import annotations

import os

import numpy as np
import torch

from .make_golden import OUT, versions
from .ref_import import build_reference_unet, load_reference
from .synth import synth_pool_u8
from .unet_ref import golden_state_dict


def main():
    ref = load_reference()
    gm = np.load(os.path.join(OUT, "model.npz"))
    out = {"versions": np.array(repr(versions()))}
    for K, name in ((2, "pool_K2"), (3, "pool_K3")):
        g = np.load(os.path.join(OUT, f"{name}.npz"))
        L, n = int(g["n_labeled"]), int(g["n_pool"])
        model = build_reference_unet(K)
        model.load_state_dict(golden_state_dict(K, gm))
        model.eval()
        loss = ref.DiceLoss(num_classes=K - 1, smooth=1e-5, do_bg=True, softmax=True)
        pool = synth_pool_u8(L + n)[L:]
        scores = []
        with torch.no_grad():
            for i in range(0, n, 8):
                logits = model(pool[i:i + 8].float().div(255).unsqueeze(1))
                for j in range(logits.shape[0]):
                    lg = logits[j:j + 1]
                    scores.append(float(loss(lg, lg.softmax(1).argmax(1))))
        out[f"scores/K{K}"] = np.array(scores, np.float32)
        print(name, out[f"scores/K{K}"][:4])
    np.savez_compressed(os.path.join(OUT, "dice.npz"), **out)


if __name__ == "__main__":
    main()
