"""Generate tests/golden/kmean.npz by running the UNMODIFIED reference KMeanSelector here (container only; needs
/root/reference).  Usage:  python -m oracle.make_golden_kmean

The golden model (tests/golden/model.npz) is loaded into the reference UNet; pools are the seeded synthetic ones of
make_golden.py (K=3: 16 labeled + 48 pool, budget 8; K=2: 32 labeled + 128 pool, budget 16; K=3 with NO labeled data:
uniform weights).  Before every call numpy's global RNG is seeded with NP_SEED - `kmeans_plusplus` draws from it
(kmean_selector.py:190-192).  Stored per case: the sorted selected names (the reference returns an unordered
`list(set(...))`, :194) and `gap/<case>` = the conditioning of the run measured with the oracle restatement
(oracle.kmean_ref.kmeans_plusplus_indices(return_gap=True)): input perturbations well below it cannot change the picks.
"""
from __future__ import annotations

import os

import numpy as np
import torch

from . import select_ref as S
from .kmean_ref import kmean_select
from .ref_import import build_reference_unet, load_reference
from .synth import make_active, synth_pool_u8
from .unet_ref import golden_state_dict

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
NP_SEED = 4242

# (key, K, n_labeled, n_pool, budget, selector kwargs)
CASES = [
    ("K3/l2/min", 3, 16, 48, 8, dict(metric="l2")),
    ("K3/cosine/min", 3, 16, 48, 8, dict(metric="cosine")),
    ("K3/l2/mean", 3, 16, 48, 8, dict(metric="l2", coreset_criteria="mean")),
    ("K3/cosine/mean/sharp2", 3, 16, 48, 8, dict(metric="cosine", coreset_criteria="mean", sharp_factor=2.0)),
    ("K3/l2/min/softmax", 3, 16, 48, 8, dict(metric="l2", softmax=True, sharp_factor=3.0)),
    ("K3/l2/cold", 3, 0, 64, 8, dict(metric="l2")),
    ("K2/l2/min", 2, 32, 128, 16, dict(metric="l2")),
    ("K2/cosine/min/sharp2", 2, 32, 128, 16, dict(metric="cosine", sharp_factor=2.0)),
    # the demo's model-free mode (entry/demo/serve.py:125-134): features come from a dict, the model is not used
    ("dict/l2/min", 3, 12, 60, 10, dict(metric="l2", loaded_feature_only=True)),
    ("dict/cosine/mean", 3, 12, 60, 10, dict(metric="cosine", coreset_criteria="mean", loaded_feature_only=True)),
]
DICT_DIM, DICT_SEED = 200, 99


def loaded_features(names) -> dict:
    """Seeded stand-in for foundation-model features: one fp32 row of DICT_DIM per case name."""
    rng = np.random.default_rng(DICT_SEED)
    base = rng.standard_normal((8, DICT_DIM))
    return {n: (base[i % 8] * (1 + 0.3 * rng.standard_normal()) + 0.5 * rng.standard_normal(DICT_DIM)).astype(np.float32)
            for i, n in enumerate(names)}


def main():
    import sklearn
    ref = load_reference()
    g = np.load(os.path.join(OUT, "model.npz"))
    torch.set_num_threads(os.cpu_count())
    out = {}
    models = {}
    for key, K, n_lab, n_pool, budget, kw in CASES:
        if K not in models:
            m = build_reference_unet(K)
            m.load_state_dict(golden_state_dict(K, g))
            models[K] = m
        ad = make_active(synth_pool_u8(n_lab + n_pool), n_lab, reference=ref)
        all_names = list(ad.get_train_dataset().image_idx) + list(ad.get_pool_dataset().image_idx)
        fd = loaded_features(all_names) if kw.get("loaded_feature_only") else None
        sel = ref.KMeanSelector(batch_size=8, num_workers=0, pin_memory=False, feature_dict=fd, **kw)
        np.random.seed(NP_SEED)
        names = sel.select_next_batch(ad, budget, None if fd else models[K], torch.device("cpu"))
        out[key] = np.array(sorted(map(str, names)))
        if fd:
            emb = np.stack([fd[n] for n in all_names])
        else:
            emb = S.pool_embeddings(golden_state_dict(K, g), synth_pool_u8(n_lab + n_pool))
        idx, gap = kmean_select(emb[n_lab:], emb[:n_lab] if n_lab else None, budget, kw.get("metric", "cosine"),
                                kw.get("coreset_criteria", "min"), kw.get("sharp_factor", 1.0), kw.get("softmax", False),
                                np.random.RandomState(NP_SEED), return_gap=True)
        assert sorted({ad.get_pool_dataset().image_idx[int(i)] for i in idx}) == out[key].tolist()  # oracle == reference
        out["gap/" + key] = np.float64(gap)
        print(key, len(names), out[key][:4], f"gap {gap:.2e}")
    out["np_seed"] = np.int64(NP_SEED)
    out["versions"] = np.array(f"sklearn={sklearn.__version__} numpy={np.__version__} torch={torch.__version__}")
    np.savez_compressed(os.path.join(OUT, "kmean.npz"), **out)


if __name__ == "__main__":
    main()
