"""tests/golden/cat_mix.npz: CoresetSelector(coreset_fusion="cat") and KMeanSelector with model + loaded features
(container only; needs /root/reference).  Usage:  python -m oracle.make_golden_cat

Both paths multiply the loaded rows by the numpy scalar np.sqrt(D1 / D2 * loaded_feature_weight)
(coreset_selector.py:166-170, kmean_selector.py:113-119).  The reference pins numpy < 2 (pyproject.toml:21), where that
product stays float32; this image has numpy 2.3.5, where it is promoted to float64 - so the reference CANNOT be run
here under its own dtype rule ("live pin unavailable with numpy 2.3.5").  What this script does instead:
  1. runs the UNMODIFIED reference here (numpy 2 semantics) and checks that the oracle restatement evaluated under the
     same NEP-50 rule reproduces it: selected sets of CoresetSelector(cat) for metric x criteria, and the feature matrix
     of KMeanSelector.get_features bit for bit - this pins everything except the dtype rule (row order, concatenation
     order, standardisation before scaling, the scale formula);
  2. freezes the selections of the SAME restatement under the legacy (numpy < 2) rule - the contract the product
     path implements - together with their conditioning gaps.
h5py is absent: the reference's module attribute is pointed at the File/Dataset stand-in of make_golden_fused.py.
"""
from __future__ import annotations

import importlib
import os
import tempfile

import numpy as np
import torch

from . import select_ref as S
from .kmean_ref import kmean_features, kmean_mix_select
from .make_golden_fused import H5Shim, write_feature_files
from .ref_import import build_reference_unet, load_reference
from .synth import make_active, synth_pool_u8
from .unet_ref import golden_state_dict

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "cat_mix.npz")
N_LAB, N_POOL, BUDGET, WEIGHT, NP_SEED = 8, 40, 7, 0.4, 777
DIMS = (200, 700)   # joint widths 712 -> 1024 columns, 1212 -> 1536 columns in the kernels


def loaded_features(names, dim) -> dict:
    rng = np.random.default_rng(11 + dim)
    base = rng.standard_normal((6, dim))
    return {n: (base[i % 6] + 0.6 * rng.standard_normal(dim)).astype(np.float32) for i, n in enumerate(names)}


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
    emb = S.pool_embeddings(sd, images)
    cpu = torch.device("cpu")
    out = {"numpy_live": np.array(np.__version__)}
    for dim in DIMS:
        feats = loaded_features(names, dim)
        loaded = np.stack([feats[n] for n in names])
        with tempfile.TemporaryDirectory() as d:
            write_feature_files(d, feats)
            for metric in ("l2", "cosine"):
                for crit in ("min", "mean"):
                    sel = ref.CoresetSelector(batch_size=8, num_workers=0, pin_memory=False, metric=metric,
                                              coreset_criteria=crit, coreset_fusion="cat", feature_path=d,
                                              loaded_feature_weight=WEIGHT)
                    live = sel.select_next_batch(make_active(images, N_LAB, reference=ref), BUDGET, model, cpu)
                    p64, g64 = S.coreset_cat_select(emb, loaded, WEIGHT, N_LAB, BUDGET, metric, crit, "nep50", True)
                    assert sorted(map(str, live)) == sorted(names[i] for i in p64), (dim, metric, crit)
                    p32, g32 = S.coreset_cat_select(emb, loaded, WEIGHT, N_LAB, BUDGET, metric, crit, "legacy", True)
                    key = f"coreset/{dim}/{metric}/{crit}"
                    out["selected/" + key] = np.array(sorted(names[i] for i in p32))
                    out["gap/" + key] = np.float64(min(g32))
                    print(key, "live(nep50) == oracle(nep50); legacy", out["selected/" + key][:4].tolist(),
                          f"gap {min(g32):.2e}", "same set as nep50" if sorted(p32) == sorted(p64) else "DIFFERS from nep50")
        # KMeanSelector.get_features, live vs restatement under the same rule
        ksel = ref.KMeanSelector(batch_size=8, num_workers=0, pin_memory=False, feature_dict=feats,
                                 loaded_feature_weight=WEIGHT, loaded_feature_only=False)
        live_feats, live_names = ksel.get_features(make_active(images, N_LAB, reference=ref).get_pool_dataset(), model, cpu)
        want = kmean_features(emb[N_LAB:], loaded[N_LAB:], WEIGHT, "nep50")
        assert list(live_names) == names[N_LAB:] and live_feats.dtype == want.dtype
        assert np.array_equal(live_feats, want), "KMeanSelector.get_features != restatement"
        print(f"kmean/{dim}: get_features live == restatement (nep50), dtype {live_feats.dtype}")
        for metric, crit, sharp in (("l2", "min", 1.0), ("cosine", "mean", 2.0)):
            seed = NP_SEED
            while True:
                idx, gap = kmean_mix_select(emb[N_LAB:], emb[:N_LAB], loaded[N_LAB:], loaded[:N_LAB], WEIGHT, BUDGET,
                                            metric, crit, sharp, False, np.random.RandomState(seed), True)
                if gap > 1e-4:
                    break
                seed += 1
            key = f"kmean/{dim}/{metric}/{crit}"
            out["selected/" + key] = np.array([names[N_LAB + int(i)] for i in idx])
            out["seed/" + key] = np.int64(seed)
            out["gap/" + key] = np.float64(gap)
            print(key, out["selected/" + key].tolist(), f"seed {seed} gap {gap:.2e}")
    np.savez_compressed(OUT, **out)


if __name__ == "__main__":
    main()
