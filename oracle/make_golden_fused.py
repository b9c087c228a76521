"""Generate tests/golden/coreset_fused.npz by running the UNMODIFIED reference CoresetSelector with externally loaded
features here (container only; needs /root/reference).  Usage:  python -m oracle.make_golden_fused

The reference reads one `{case}.h5` per image through h5py (coreset_selector.py:110-119), which is not installed in
this image.  Only for this script (and the matching reference test) the module attribute `h5py` inside the reference's
`activelearning.coreset_selector` is pointed at a 20-line stand-in whose `File(path)["feature"][:]` returns an array
saved with numpy under that file name; everything downstream of the read - `pairwise_distances`, the global `/sum`
normalisations, the weighted add, `kcenter_greedy`, the cold-start `kmeans_plusplus` - is the reference's own code.

Golden model K=3, pool = 8 labeled + 40 pool synthetic images, loaded features = `loaded_features(names)` (seeded,
200 columns), loaded_feature_weight 0.4, coreset_fusion "add".  Stored: sorted selected names per (metric, criteria),
the conditioning gap of each run (oracle.select_ref.kcenter_fused_add_stream), and the cold-start (no labeled data)
k-means++ picks in order with numpy's global RNG seeded with NP_SEED.
"""
from __future__ import annotations

import importlib
import os
import tempfile

import numpy as np
import torch

from . import select_ref as S
from .kmean_ref import kmeans_plusplus_indices
from .ref_import import build_reference_unet, load_reference
from .synth import make_active, synth_pool_u8
from .unet_ref import golden_state_dict

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "coreset_fused.npz")
N_LAB, N_POOL, BUDGET, WEIGHT, NP_SEED, DIM = 8, 40, 7, 0.4, 2024, 200


def loaded_features(names) -> dict:
    rng = np.random.default_rng(5)
    base = rng.standard_normal((6, DIM))
    return {n: (base[i % 6] + 0.6 * rng.standard_normal(DIM)).astype(np.float32) for i, n in enumerate(names)}


class _Dataset:
    def __init__(self, a):
        self.a = a

    def __getitem__(self, k):
        return self.a[k]


class _File:
    def __init__(self, path, mode="r"):
        with open(path, "rb") as f:
            self.d = {"feature": _Dataset(np.load(f))}

    def __enter__(self):
        return self.d

    def __exit__(self, *a):
        return False


class H5Shim:
    """Stands in for the `h5py` module inside the reference's coreset_selector (File + Dataset only)."""
    File = _File
    Dataset = _Dataset


def write_feature_files(folder: str, feats: dict) -> None:
    for name, v in feats.items():
        with open(os.path.join(folder, f"{name}.h5"), "wb") as f:
            np.save(f, v)


def main():
    ref = load_reference()
    cs = importlib.import_module("activelearning.coreset_selector")
    cs.h5py = H5Shim
    g = np.load(os.path.join(os.path.dirname(OUT), "model.npz"))
    sd = golden_state_dict(3, g)
    model = build_reference_unet(3)
    model.load_state_dict(sd)
    images = synth_pool_u8(N_LAB + N_POOL)
    ad = make_active(images, N_LAB, reference=ref)
    names = list(ad.get_train_dataset().image_idx) + list(ad.get_pool_dataset().image_idx)
    feats = loaded_features(names)
    loaded = np.stack([feats[n] for n in names])
    emb = S.pool_embeddings(sd, images)
    out = {}
    with tempfile.TemporaryDirectory() as d:
        write_feature_files(d, feats)
        kw = dict(batch_size=8, num_workers=0, pin_memory=False, coreset_fusion="add", feature_path=d,
                  loaded_feature_weight=WEIGHT)
        for metric in ("l2", "cosine"):
            for crit in ("min", "mean"):
                sel = ref.CoresetSelector(metric=metric, coreset_criteria=crit, **kw)
                want = sel.select_next_batch(make_active(images, N_LAB, reference=ref), BUDGET, model, torch.device("cpu"))
                picks, gaps = S.kcenter_fused_add_stream(emb, loaded, WEIGHT, N_LAB, BUDGET, metric, crit, True)
                assert sorted(map(str, want)) == sorted(names[i] for i in picks)
                out[f"selected/{metric}/{crit}"] = np.array(sorted(map(str, want)))
                out[f"gap/{metric}/{crit}"] = np.float64(min(gaps))
                print(metric, crit, sorted(map(str, want))[:4], f"gap {min(gaps):.2e}")
        # cold start with loaded features: kmeans_plusplus over the loaded rows (coreset_selector.py:202-210)
        ad0 = make_active(images, 0, reference=ref)
        sel = ref.CoresetSelector(metric="l2", **kw)
        seed = NP_SEED
        while True:  # a well-conditioned seed (see oracle/make_golden_badge.py)
            idx, gap = kmeans_plusplus_indices(loaded, BUDGET, None, np.random.RandomState(seed), return_gap=True)
            if gap > 1e-4:
                break
            seed += 1
        np.random.seed(seed)
        cold = sel.select_next_batch(ad0, BUDGET, model, torch.device("cpu"))
        assert list(map(str, cold)) == [names[i] for i in idx]
        out["selected/cold"] = np.array(list(map(str, cold)))
        out["seed/cold"] = np.int64(seed)
        print("cold", list(map(str, cold)), f"seed {seed} gap {gap:.2e}")
    np.savez_compressed(OUT, **out)


if __name__ == "__main__":
    main()
