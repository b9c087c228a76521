"""Generate tests/golden/* by running the UNMODIFIED reference here (container
only; needs /root/reference).  Usage:  python -m oracle.make_golden

Files written (all small; inputs are regenerated from seeds, not stored):
  model.npz      BN tensors + K=2 / K=3 heads of the golden model, conv digest
  pool_K2.npz    BUSI-shape (K=2): 32 labeled + 128 pool images: reference
                 scores (entropy/margin/confidence), embeddings, selected names
                 of the reference selectors (budget 16), probe logits
  pool_K3.npz    FUGC-shape (K=3): same for 16 labeled + 48 pool images
  mc_K3.npz      composed MC-dropout oracle (T=4, p=0.1) on 8 pool images
Versions of torch / scikit-learn / numpy that produced them are recorded.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

from . import select_ref as S
from .ref_import import build_reference_unet, load_reference
from .synth import SEED, make_active, synth_pool_u8
from .unet_ref import GOLDEN_BN_KEYS, build_calibrated_state_dict, conv_weight_digest

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
MC_SEED = 2025


def versions():
    import sklearn
    return dict(torch=torch.__version__, sklearn=sklearn.__version__, numpy=np.__version__)


def run_reference_pool(K, n_labeled, n_pool, budget, calib):
    ref = load_reference()
    sd = build_calibrated_state_dict(K, calib)
    model = build_reference_unet(K)
    model.load_state_dict(sd)
    images = synth_pool_u8(n_labeled + n_pool)
    out = {}
    cpu = torch.device("cpu")
    kw = dict(batch_size=8, num_workers=0, pin_memory=False)
    for mode, cls in (("entropy", ref.EntropySelector), ("margin", ref.MarginSelector),
                      ("confidence", ref.ConfidenceSelector)):
        ad = make_active(images, n_labeled, reference=ref)
        sel = cls(**kw)
        score_list, names = sel.cal_scores(ad, model, cpu)
        out[f"scores/{mode}"] = torch.stack(score_list).numpy()
        ad = make_active(images, n_labeled, reference=ref)
        out[f"selected/{mode}"] = np.array(sel.select_next_batch(ad, budget, model, cpu))
        assert list(names) == ad.get_pool_dataset().image_idx
    for metric in ("l2", "cosine"):
        ad = make_active(images, n_labeled, reference=ref)
        sel = ref.CoresetSelector(metric=metric, **kw)
        core, all_list, _lf, feats, dist = sel.cal_scores(ad, model, cpu)
        out["embeddings"] = feats.astype(np.float32)
        ids = ref.kcenter_greedy(dist, len(all_list), budget, np.arange(len(core)), "min")
        out[f"selected/coreset-{metric}"] = np.array(sorted(all_list[ids].tolist()))
        ad = make_active(images, n_labeled, reference=ref)
        names = sel.select_next_batch(ad, budget, model, cpu)
        assert sorted(map(str, names)) == out[f"selected/coreset-{metric}"].tolist()
    # probe logits: first 2 pool images
    x = images[n_labeled:n_labeled + 2].float().div(255).unsqueeze(1)
    model.eval()
    with torch.no_grad():
        out["probe_logits"] = model(x).numpy()
    out["n_labeled"] = np.int64(n_labeled)
    out["n_pool"] = np.int64(n_pool)
    out["budget"] = np.int64(budget)
    out["versions"] = np.array(repr(versions()))
    return sd, out


def main():
    if not os.path.isdir("/root/reference"):
        sys.exit("needs /root/reference")
    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(os.cpu_count())
    calib = synth_pool_u8(64, start=900_000)  # calibration batch, disjoint from pools

    sd2, g2 = run_reference_pool(2, 32, 128, 16, calib)
    np.savez_compressed(os.path.join(OUT, "pool_K2.npz"), **g2)
    sd3, g3 = run_reference_pool(3, 16, 48, 8, calib)
    np.savez_compressed(os.path.join(OUT, "pool_K3.npz"), **g3)

    model = {"conv_digest": np.array(conv_weight_digest(sd3)), "versions": np.array(repr(versions()))}
    assert conv_weight_digest(sd2) == conv_weight_digest(sd3)
    for k, v in sd3.items():
        if k.endswith(GOLDEN_BN_KEYS):
            assert torch.equal(v, sd2[k])
            model["bn/" + k] = v.numpy()
    for K, sd in ((2, sd2), (3, sd3)):
        model[f"head{K}/weight"] = sd["decoder.seg_output.weight"].numpy()
        model[f"head{K}/bias"] = sd["decoder.seg_output.bias"].numpy()
    np.savez_compressed(os.path.join(OUT, "model.npz"), **model)

    # composed MC-dropout oracle (reference UNet module with Dropout2d replaced
    # by explicit masks == oracle.unet_ref with masks; pinned in tests)
    images = synth_pool_u8(8, start=16)
    pos = np.arange(8)
    mc = {}
    for mode in ("entropy", "margin", "confidence"):
        mc[f"scores/{mode}"] = S.pool_scores_mc(sd3, images, pos, 4, 0.1, MC_SEED, mode)
    mc["T"] = np.int64(4)
    mc["p"] = np.float64(0.1)
    mc["seed"] = np.int64(MC_SEED)
    mc["start"] = np.int64(16)
    np.savez_compressed(os.path.join(OUT, "mc_K3.npz"), **mc)
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main()
