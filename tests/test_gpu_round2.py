"""GPU parity of the rows added in round 2: dice-uncertainty (kernel 3, third mode), coreset_fusion="cat", KMeanSelector
with model + loaded features, feature_path files, and the non-finite guards."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

DEV = "cuda:0"
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


@pytest.fixture(scope="module")
def k3(golden):
    import b200al
    from oracle.unet_ref import OracleUNet, golden_state_dict
    sd = golden_state_dict(3, golden["model"])
    eng = b200al.Engine(DEV, 3, max_samples=32)
    eng.load_model(sd)
    yield dict(sd=sd, eng=eng, model=OracleUNet(sd))
    eng.close()


@pytest.mark.parametrize("K", [2, 3])
def test_dice_scores_match_reference_diceloss(golden, K):
    """scores frozen with the reference's own DiceLoss class on reference logits (oracle/make_golden_dice.py)."""
    import b200al
    from oracle.synth import synth_pool_u8
    from oracle.unet_ref import golden_state_dict
    g = golden[f"pool_K{K}"]
    L, n = int(g["n_labeled"]), int(g["n_pool"])
    want = np.load(os.path.join(GOLDEN, "dice.npz"))[f"scores/K{K}"]
    eng = b200al.Engine(DEV, K, max_samples=32)
    eng.load_model(golden_state_dict(K, golden["model"]))
    got = eng.scores(synth_pool_u8(L + n)[L:].to(DEV), "dice", smooth=1e-5).cpu().numpy()
    rel = np.abs(got - want) / np.abs(want)
    assert rel.max() < 1e-3, f"max rel err {rel.max():.3e}"
    eng.close()


def test_dice_kernel_on_logits_and_mc(k3):
    from oracle import select_ref as S
    eng = k3["eng"]
    g = torch.Generator().manual_seed(3)
    for K, T, hw in ((3, 1, 64), (2, 1, 128), (4, 5, 64), (3, 8, 256)):
        lg = torch.randn(3, T, K, hw, hw, generator=g) * 2
        got = eng.uncertainty_from_logits(lg.to(DEV), "dice", smooth=1e-5).cpu().numpy()
        want = S.dice_scores_from_prob(lg.softmax(2).mean(1)).numpy()
        np.testing.assert_allclose(got, want, rtol=1e-4)
    # through the selector, with MC repeats: composed oracle (reference UNet + masks, mean prob, DiceLoss formula)
    from oracle.synth import synth_pool_u8
    imgs = synth_pool_u8(4, start=16)
    got = eng.scores(imgs.to(DEV), "dice", T=4, drop_p=0.1, seed=5, base_pos=0, smooth=1e-5).cpu().numpy()
    want = S.pool_scores_mc(k3["sd"], imgs, np.arange(4), 4, 0.1, 5, "dice")
    assert (np.abs(got - want) / np.abs(want)).max() < 1e-3


def test_dice_selector_picks_oracle_order(k3, golden):
    import b200al
    from oracle import select_ref as S
    from oracle.synth import make_active, synth_pool_u8
    g = golden["pool_K3"]
    L, n = int(g["n_labeled"]), int(g["n_pool"])
    images = synth_pool_u8(L + n)
    ad = make_active(images, L)
    want_scores = np.load(os.path.join(GOLDEN, "dice.npz"))["scores/K3"]
    sel = b200al.B200DiceUncertaintySelector(batch_size=8, num_workers=0, pin_memory=False, chunk_samples=32)
    got = sel.select_next_batch(ad, 8, k3["model"], torch.device(DEV))
    want = [ad.get_pool_dataset().image_idx[i] for i in S.rank_desc_stable(want_scores, 8)]
    scores, _ = sel.cal_scores(ad, k3["model"], torch.device(DEV))
    err = np.abs(scores.cpu().numpy() - want_scores).max()
    assert S.boundary_gap(want_scores, 8) > 10 * err, "golden cut is a near-tie relative to the measured error"
    assert got == want


def _cat_env(dim):
    from oracle.make_golden_cat import N_LAB, N_POOL, loaded_features
    from oracle.synth import case_name, synth_pool_u8
    names = [case_name(i) for i in range(N_LAB + N_POOL)]
    return names, loaded_features(names, dim), synth_pool_u8(N_LAB + N_POOL)


@pytest.mark.parametrize("dim", [200, 700])
def test_coreset_cat_fusion_reproduces_frozen_selection(k3, dim, tmp_path):
    """coreset_fusion="cat" (coreset_selector.py:158-175) under the reference's numpy < 2 dtype rule; the frozen
    sets equal what the live reference selected here under numpy 2 (oracle/make_golden_cat.py)."""
    import b200al
    from oracle.make_golden_cat import BUDGET, N_LAB, WEIGHT
    from oracle.synth import make_active
    gc = np.load(os.path.join(GOLDEN, "cat_mix.npz"))
    names, fd, images = _cat_env(dim)
    for n, v in fd.items():            # feature_path route: one file per case (.npy where h5py is absent)
        np.save(tmp_path / f"{n}.npy", v)
    for metric in ("l2", "cosine"):
        for crit in ("min", "mean"):
            key = f"coreset/{dim}/{metric}/{crit}"
            kw = dict(batch_size=8, num_workers=0, pin_memory=False, metric=metric, coreset_criteria=crit,
                      coreset_fusion="cat", loaded_feature_weight=WEIGHT, chunk_samples=32)
            for src in (dict(feature_dict=fd), dict(feature_path=str(tmp_path))):
                sel = b200al.B200CoresetSelector(**kw, **src)
                got = sel.select_next_batch(make_active(images, N_LAB), BUDGET, k3["model"], torch.device(DEV))
                assert float(gc["gap/" + key]) > 1e-4
                assert sorted(got) == gc["selected/" + key].tolist(), key


@pytest.mark.parametrize("dim", [200, 700])
def test_kmean_model_plus_loaded_features(k3, dim):
    """KMeanSelector with model AND loaded features (kmean_selector.py:106-121), numpy < 2 rule; pick order frozen
    from the oracle restatement whose feature matrix is pinned bit for bit against the live reference."""
    import b200al
    from oracle.make_golden_cat import BUDGET, N_LAB, WEIGHT
    from oracle.synth import make_active
    gc = np.load(os.path.join(GOLDEN, "cat_mix.npz"))
    names, fd, images = _cat_env(dim)
    for metric, crit, sharp in (("l2", "min", 1.0), ("cosine", "mean", 2.0)):
        key = f"kmean/{dim}/{metric}/{crit}"
        sel = b200al.B200KMeanSelector(batch_size=8, num_workers=0, pin_memory=False, metric=metric,
                                       coreset_criteria=crit, sharp_factor=sharp, feature_dict=fd,
                                       loaded_feature_weight=WEIGHT, chunk_samples=32)
        np.random.seed(int(gc["seed/" + key]))
        got = sel.select_next_batch(make_active(images, N_LAB), BUDGET, k3["model"], torch.device(DEV))
        assert got == gc["selected/" + key].tolist(), key


def test_wide_rows_kcenter_and_kmeanspp_exact():
    """1536- and 2048-column rows (NV = 12 / 16 kernels) against the streaming oracle / the k-means++ restatement."""
    import b200al
    from oracle import select_ref as S
    from oracle.kmean_ref import kmeans_plusplus_indices
    eng = b200al.Engine(DEV, 2, max_samples=1)
    rng = np.random.default_rng(5)
    for D in (1536, 2048):
        feats = (rng.standard_normal((600, D)) * 0.3 + 0.1).astype(np.float32)
        for metric in ("l2", "cosine"):
            want = S.kcenter_greedy_stream(feats, 9, 40, metric)
            assert eng.kcenter(torch.from_numpy(feats).to(DEV), 9, 40, metric).cpu().tolist() == want, (D, metric)
        k = 12
        trials = 2 + int(np.log(k))
        for seed in range(3):
            idx, gap = kmeans_plusplus_indices(feats, k, None, np.random.RandomState(seed), return_gap=True)
            if gap < 1e-5:
                continue
            rs = np.random.RandomState(seed)
            w = np.ones(600, np.float32)
            first = int(rs.choice(600, p=w / w.sum()))
            uni = np.stack([rs.uniform(size=trials) for _ in range(k - 1)])
            got = eng.kmeans_pp(torch.from_numpy(feats).to(DEV), torch.from_numpy(w), first, torch.from_numpy(uni), k, trials)
            assert got.cpu().tolist() == idx.tolist(), (D, seed)
    eng.close()


def test_non_finite_embeddings_raise(k3):
    """A diverged model must not return an arbitrary selection (the argmax kernels never pick a NaN row)."""
    import b200al
    from oracle.synth import make_active, synth_pool_u8
    from oracle.unet_ref import OracleUNet
    sd = {k: v.clone() for k, v in k3["sd"].items()}
    sd["encoder.levels.4.1.all.2.weight"][:] = float("nan")
    model = OracleUNet(sd)
    ad = make_active(synth_pool_u8(24), 4)
    for sel in (b200al.B200CoresetSelector(8, 0, False, metric="l2", chunk_samples=32),
                b200al.B200HybridSelector(8, 0, False, chunk_samples=32),
                b200al.B200KMeanSelector(8, 0, False, metric="l2", chunk_samples=32)):
        with pytest.raises(RuntimeError, match="non-finite"):
            sel.select_next_batch(ad, 4, model, torch.device(DEV))
    eng = k3["eng"]
    eng.load_model(k3["sd"])  # other tests share the engine
