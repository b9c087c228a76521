"""KMeanSelector on the GPU: row standardisation, pool-to-labeled weights and scikit-learn's
weighted k-means++ seeding as CUDA kernels, against the oracle restatement (oracle/kmean_ref.py, pinned to the
reference's KMeanSelector), scikit-learn's own `kmeans_plusplus` (installed in this image) and the frozen reference
selections in tests/golden/kmean.npz."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _engine():
    import b200al
    return b200al.Engine(DEV, 2, max_samples=1)


def test_standardize_and_centre_distances_match_numpy_and_sklearn():
    from sklearn.metrics import pairwise_distances
    from oracle.kmean_ref import standardize_rows
    rng = np.random.default_rng(1)
    # rows with spread 0.5..3 around offsets ~N(0,1): numpy's fp32 row mean is exact to ~1e-7, the kernel's comes from fp64
    feats = (rng.standard_normal((300, 512)) * (rng.random((300, 1)) * 2.5 + 0.5) + rng.standard_normal((300, 1))).astype(np.float32)
    eng = _engine()
    try:
        got = eng.standardize_rows(torch.from_numpy(feats).to(DEV)).cpu().numpy()
        want = standardize_rows(feats)
        assert np.abs(got - want).max() < 5e-6
        X, Xl = want[40:], want[:40]
        for metric in ("l2", "cosine"):
            dist = pairwise_distances(X, Xl, metric=metric)  # what the reference calls (kmean_selector.py:143-145)
            for criteria, red in (("min", dist.min(axis=1)), ("mean", dist.mean(axis=1))):
                w = eng.centre_distances(torch.from_numpy(X).to(DEV), torch.from_numpy(Xl).to(DEV), metric, criteria)
                np.testing.assert_allclose(w.cpu().numpy(), red, rtol=2e-6, atol=2e-7)
    finally:
        eng.close()


def test_kmeans_pp_kernels_pick_sklearns_indices():
    """Index-exact against the oracle on every case, and against scikit-learn itself (global numpy RNG seeded, same
    draw order) wherever the run is well conditioned: sklearn sums its potentials with an fp32 BLAS dot of
    CPU-specific order (~1e-6 relative), so when a uniform draw lies closer than that to a cumulative-weight boundary
    (large n, many draws) sklearn's own picks depend on the host CPU and are not a reference."""
    from sklearn.cluster import kmeans_plusplus
    from oracle.kmean_ref import kmeans_plusplus_indices, standardize_rows
    rng = np.random.default_rng(7)
    eng = _engine()
    vs_sklearn = 0
    try:
        for trial, (n, k) in enumerate([(64, 5), (300, 24), (1000, 40), (5000, 64), (9000, 1), (12, 12)]):
            X = standardize_rows(rng.standard_normal((n, 512)).astype(np.float32) * rng.random((n, 1)).astype(np.float32)
                                 + rng.standard_normal((1, 512)).astype(np.float32))
            w = rng.random(n).astype(np.float32)
            w /= w.sum()
            for weights in (w, None):
                np.random.seed(100 + trial)
                _, want = kmeans_plusplus(X, k, sample_weight=weights)
                oracle, gap = kmeans_plusplus_indices(X, k, weights, np.random.RandomState(100 + trial), return_gap=True)
                rs = np.random.RandomState(100 + trial)
                wk = np.ones(n, np.float32) if weights is None else weights
                trials = 2 + int(np.log(k))
                first = int(rs.choice(n, p=wk / wk.sum()))
                u = np.stack([rs.uniform(size=trials) for _ in range(k - 1)]) if k > 1 else np.zeros((0, trials))
                got = eng.kmeans_pp(torch.from_numpy(X).to(DEV), torch.from_numpy(wk), first, torch.from_numpy(u), k,
                                    trials).cpu().numpy()
                assert np.array_equal(got, oracle), (n, k, weights is None)
                if gap > 5e-6:
                    vs_sklearn += 1
                    assert np.array_equal(got, want), (n, k, weights is None, gap)
        assert vs_sklearn >= 8
    finally:
        eng.close()


def test_kmean_selector_matches_reference_golden(golden):
    """The drop-in selector end to end (embeddings from the tcgen05 encoder) vs the reference KMeanSelector's frozen
    selections, all ten cases (l2 / cosine, min / mean, sharp_factor, softmax, no labeled data, feature_dict mode)."""
    import b200al
    from oracle.make_golden_kmean import CASES, loaded_features
    from oracle.synth import case_name, make_active, synth_pool_u8
    from oracle.unet_ref import OracleUNet, golden_state_dict
    gk = np.load(os.path.join(os.path.dirname(__file__), "golden", "kmean.npz"))
    models = {K: OracleUNet(golden_state_dict(K, golden["model"])) for K in (2, 3)}
    for key, K, n_lab, n_pool, budget, kw in CASES:
        ad = make_active(synth_pool_u8(n_lab + n_pool), n_lab)
        # the demo's model-free mode: features per case name from a dict (200 columns -> padded to 256), no model
        fd = loaded_features([case_name(i) for i in range(n_lab + n_pool)]) if kw.get("loaded_feature_only") else None
        sel = b200al.B200KMeanSelector(batch_size=8, num_workers=0, pin_memory=False, feature_dict=fd, **kw)
        np.random.seed(int(gk["np_seed"]))
        names = sel.select_next_batch(ad, budget, None if fd else models[K], torch.device(DEV))
        assert len(names) == len(set(names)) and set(names) <= set(ad.get_pool_dataset().image_idx)
        assert sorted(names) == gk[key].tolist(), key
    # the numpy stream advanced exactly as far as sklearn advances it: 1 choice + (k-1) * trials uniforms
    np.random.seed(5)
    ad = make_active(synth_pool_u8(16 + 48), 16)
    b200al.B200KMeanSelector(batch_size=8, num_workers=0, metric="l2").select_next_batch(ad, 8, models[3], torch.device(DEV))
    after = np.random.random_sample()
    np.random.seed(5)
    np.random.random_sample(1 + 7 * (2 + int(np.log(8))))
    assert after == np.random.random_sample()


def test_kmean_selector_contract_and_registry():
    import b200al
    with pytest.raises(NotImplementedError):
        b200al.B200KMeanSelector(batch_size=1, num_workers=0, metric="l1")
    # round 2: feature_path (.h5 / .npy / .npz per case) and the model + loaded mix are supported
    # (tests/test_gpu_round2.py); what stays refused is loaded_feature_only without any loaded features
    assert str(b200al.B200KMeanSelector(batch_size=1, num_workers=0, feature_path="/tmp/x").feature_path) == "/tmp/x"
    assert b200al.B200KMeanSelector(batch_size=1, num_workers=0, feature_dict={"a": [1.0, 2.0]}).has_loaded
    with pytest.raises(ValueError):
        b200al.B200KMeanSelector(batch_size=1, num_workers=0, loaded_feature_only=True)
    with pytest.raises(RuntimeError):
        b200al.B200KMeanSelector(batch_size=1, num_workers=0, coreset_criteria="median")
    s = b200al.SELECTORS["b200-kmean-l2"](batch_size=2, num_workers=0, pin_memory=False)
    assert isinstance(s, b200al.B200KMeanSelector) and s.metric == "l2"
    eng = _engine()
    try:
        X = torch.randn(10, 512, device=DEV)
        with pytest.raises(RuntimeError):  # more clusters than samples: sklearn's ValueError, surfaced by the ABI too
            eng.kmeans_pp(X, torch.ones(10), 0, torch.zeros(10 * 4, dtype=torch.float64), 11, 4)
    finally:
        eng.close()
