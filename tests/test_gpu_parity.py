"""Parity of the CUDA path (through the C ABI / selectors) with the oracle and
the golden vectors frozen from the reference.  Tolerances are the north
star's: selected set identical, scores <= 1e-3 relative, logits <= 2e-2 abs."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

SCORE_RTOL = 1e-3
LOGIT_ATOL = 2e-2
DEV = "cuda:0"


@pytest.fixture(scope="module", params=["mixed", "bf16x3"])
def env(golden, request):
    """Engines in the default per-layer mix and in bf16x3 throughout: every engine-level parity test runs on both."""
    import b200al
    from oracle.synth import synth_pool_u8
    from oracle.unet_ref import OracleUNet, golden_state_dict
    e = {"precision": request.param}
    for K, name in ((2, "pool_K2"), (3, "pool_K3")):
        g = golden[name]
        n = int(g["n_labeled"]) + int(g["n_pool"])
        sd = golden_state_dict(K, golden["model"])
        eng = b200al.Engine(DEV, K, max_samples=32, precision=request.param)
        eng.load_model(sd)
        e[K] = dict(g=g, sd=sd, eng=eng, images=synth_pool_u8(n), model=OracleUNet(sd),
                    L=int(g["n_labeled"]), k=int(g["budget"]))
    yield e
    for K in (2, 3):
        e[K]["eng"].close()


def _names(idx, base=0):
    from oracle.synth import case_name
    return [case_name(base + int(i)) for i in idx]


@pytest.mark.parametrize("K", [2, 3])
@pytest.mark.parametrize("mode", ["entropy", "margin", "confidence"])
def test_scores_match_reference(env, K, mode):
    e = env[K]
    pool = e["images"][e["L"]:].to(DEV)
    got = e["eng"].scores(pool, mode).cpu().numpy()
    want = e["g"][f"scores/{mode}"]
    rel = np.abs(got - want) / np.abs(want)
    assert rel.max() < SCORE_RTOL, f"max rel err {rel.max():.3e}"


@pytest.mark.parametrize("K", [2, 3])
def test_logits_match_reference(env, K):
    e = env[K]
    pool = e["images"][e["L"]:e["L"] + 2].to(DEV)
    _s, logits = e["eng"].scores(pool, "entropy", return_logits=True)
    err = np.abs(logits.cpu().numpy() - e["g"]["probe_logits"]).max()
    assert err < LOGIT_ATOL, f"max abs logit err {err:.3e}"


@pytest.mark.parametrize("K", [2, 3])
@pytest.mark.parametrize("name", ["entropy", "margin", "confidence"])
def test_uncertainty_selectors_pick_reference_set(env, K, name):
    import b200al
    from oracle import select_ref as S
    from oracle.synth import make_active
    e = env[K]
    cls = {"entropy": b200al.B200EntropySelector, "margin": b200al.B200MarginSelector,
           "confidence": b200al.B200ConfidenceSelector}[name]
    sel = cls(batch_size=8, num_workers=0, pin_memory=True, chunk_samples=32)
    ad = make_active(e["images"], e["L"])
    got = sel.select_next_batch(ad, e["k"], e["model"], torch.device(DEV))
    want = e["g"][f"selected/{name}"].tolist()
    # the golden cut is not a near-tie: its gap is at least 10x the error the GPU scores actually show
    sc = e["g"][f"scores/{name}"]
    gpu_scores, _ = sel.cal_scores(make_active(e["images"], e["L"]), e["model"], torch.device(DEV))
    err = np.abs(gpu_scores.cpu().numpy() - sc).max()
    assert S.boundary_gap(sc, e["k"]) > 10 * err, (S.boundary_gap(sc, e["k"]), err)
    assert got == want            # same names, same order (descending score)
    assert len(set(got)) == len(got) and all(n in ad.get_pool_dataset().image_idx for n in got)
    ad.extend_train_set(got)      # what the trainer does next (active_dataset.py:41-44)
    assert ad.get_size() == (e["L"] + e["k"], len(e["images"]) - e["L"] - e["k"])


@pytest.mark.parametrize("K", [2, 3])
def test_embeddings_match_reference(env, K):
    e = env[K]
    got = e["eng"].embed(e["images"].to(DEV)).cpu().numpy()
    want = e["g"]["embeddings"]
    assert got.shape == want.shape
    err = np.abs(got - want).max() / np.abs(want).max()
    assert err < 1e-3, f"embedding err {err:.3e}"


@pytest.mark.parametrize("K", [2, 3])
@pytest.mark.parametrize("metric", ["l2", "cosine"])
def test_kcenter_kernel_exact_on_reference_embeddings(env, K, metric):
    """Stage-isolated kernel 4: same embeddings in -> identical picks, in order."""
    from oracle import select_ref as S
    e = env[K]
    feats = e["g"]["embeddings"]
    want = S.kcenter_greedy_stream(feats, e["L"], e["k"], metric)
    got, vals = e["eng"].kcenter(torch.from_numpy(feats).to(DEV), e["L"], e["k"], metric, return_vals=True)
    assert got.cpu().tolist() == want
    assert sorted(_names(want)) == e["g"][f"selected/coreset-{metric}"].tolist()


@pytest.mark.parametrize("K", [2, 3])
@pytest.mark.parametrize("metric", ["l2", "cosine"])
def test_coreset_selector_end_to_end(env, K, metric):
    import b200al
    from oracle import select_ref as S
    from oracle.synth import make_active
    e = env[K]
    sel = b200al.B200CoresetSelector(batch_size=8, num_workers=0, pin_memory=True, metric=metric, chunk_samples=32)
    ad = make_active(e["images"], e["L"])
    got = sel.select_next_batch(ad, e["k"], e["model"], torch.device(DEV))
    assert len(got) == e["k"] and len(set(got)) == e["k"]
    if sorted(got) == e["g"][f"selected/coreset-{metric}"].tolist():
        return
    # A different set is acceptable only if the reference's own greedy run had a near-tie
    # (top-2 gap below what a 1e-4 relative embedding error can move a distance by).
    _p, gaps = S.kcenter_greedy_stream(e["g"]["embeddings"], e["L"], e["k"], metric, return_gaps=True)
    bound = 2 * np.sqrt(512) * 1e-4 * np.abs(e["g"]["embeddings"]).max()
    assert min(gaps) < bound, f"picked a different set although the golden run has no near-tie (min gap {min(gaps):.2e})"
    pytest.xfail(f"golden k-center run has a near-tie (min top-2 gap {min(gaps):.2e} < {bound:.2e}): a different set "
                 "is within the stated embedding tolerance, reported as an expected failure, not hidden as a skip")


def test_kcenter_random_features_exact():
    import b200al
    from oracle import select_ref as S
    eng = b200al.Engine(DEV, 2, max_samples=1)
    rng = np.random.default_rng(7)
    for D, n, L, k in ((512, 4000, 37, 120), (128, 1000, 1, 50), (1024, 300, 5, 295)):
        feats = (rng.standard_normal((n, D)) * 0.3 + 0.1).astype(np.float32)
        for metric in ("l2", "cosine"):
            want = S.kcenter_greedy_stream(feats, L, k, metric)
            got = eng.kcenter(torch.from_numpy(feats).to(DEV), L, k, metric).cpu().tolist()
            assert got == want, (D, n, metric)
    # duplicates: exact ties resolve to the lowest row; a pool with fewer rows than the budget stops early.
    # (Expectation uses the kernel's difference-form distance: sklearn's xx+yy-2xy form leaves ~1e-7
    # rounding noise between bit-identical rows, so the reference's order among duplicates is noise.)
    feats = np.repeat(rng.standard_normal((6, 128)).astype(np.float32), 3, axis=0)
    want = S.kcenter_greedy_stream(feats, 1, 17, "l2", row_fn=S.l2_rows_direct)
    assert eng.kcenter(torch.from_numpy(feats).to(DEV), 1, 50, "l2").cpu().tolist() == want
    with pytest.raises(ValueError):
        eng.kcenter(torch.from_numpy(feats).to(DEV), 0, 3, "l2")
    eng.close()


def test_kcenter_mean_criteria_matches_oracle():
    """coreset_criteria='mean' (coreset_selector.py:35-36): the running-sum kernel picks the oracle's rows
    whenever the oracle's own top-2 margin is above fp32 rounding of the mean."""
    import b200al
    from oracle import select_ref as S
    eng = b200al.Engine(DEV, 2, max_samples=1)
    rng = np.random.default_rng(21)
    checked = 0
    for D, n, L, k in ((512, 3000, 40, 60), (256, 800, 3, 30)):
        feats = (rng.standard_normal((n, D)) * 0.3 + 0.1).astype(np.float32)
        for metric in ("l2", "cosine"):
            want, gaps = S.kcenter_greedy_mean_stream(feats, L, k, metric, return_gaps=True)
            got, vals = eng.kcenter(torch.from_numpy(feats).to(DEV), L, k, metric, return_vals=True, criteria="mean")
            got = got.cpu().tolist()
            # compare up to the first step whose margin is within rounding noise (after it the cascades may differ)
            tol = 4e-7 * float(np.abs(vals.cpu().numpy()).max())
            upto = next((i for i, g in enumerate(gaps) if g <= tol), k)
            assert got[:upto] == want[:upto], (D, metric, upto)
            checked += upto
            assert len(set(got)) == k and min(got) >= L
    # the four runs above have no step that close to a tie (smallest oracle margin 9.7e-6): every pick is compared
    assert checked == 2 * 60 + 2 * 30
    with pytest.raises(RuntimeError):
        eng.kcenter(torch.from_numpy(feats).to(DEV), L, 3, "l2", criteria="median")
    eng.close()


def test_topk_exact_with_ties():
    import b200al
    from oracle.select_ref import rank_desc_stable
    eng = b200al.Engine(DEV, 2, max_samples=1)
    rng = np.random.default_rng(1)
    cases = [
        (rng.integers(0, 30, 10007).astype(np.float32), 100),
        (rng.standard_normal(100_000).astype(np.float32), 5000),
        (np.zeros(513, np.float32), 40),
        (-np.abs(rng.standard_normal(77)).astype(np.float32), 77),
        (rng.standard_normal(5).astype(np.float32), 9),
        (np.array([3.0], np.float32), 1),
    ]
    for sc, k in cases:
        vals, pos = eng.topk(torch.from_numpy(sc).to(DEV), k, base_pos=1000)
        want = rank_desc_stable(sc, k)
        assert (pos.cpu().numpy() - 1000).tolist() == want.tolist()
        np.testing.assert_array_equal(vals.cpu().numpy(), sc[want])
    v, p = eng.topk(torch.zeros(0, device=DEV), 5)
    assert v.numel() == 0 and p.numel() == 0
    eng.close()


@pytest.mark.parametrize("K", [2, 3])
def test_uncertainty_kernel_on_reference_logits(env, K):
    """Kernel 3 alone on the reference's own logits."""
    from oracle import select_ref as S
    e = env[K]
    logits = torch.from_numpy(e["g"]["probe_logits"])
    for mode in ("entropy", "margin", "confidence"):
        got = e["eng"].uncertainty_from_logits(logits.to(DEV), mode).cpu().numpy()
        want = S.SCORE_FNS[mode](logits).numpy()
        np.testing.assert_allclose(got, want, rtol=2e-5)
    # T > 1: mean of softmax over repeats, then the score
    g = torch.Generator().manual_seed(0)
    lg = torch.randn(3, 5, K, 64, 64, generator=g) * 2
    for mode in ("entropy", "margin", "confidence"):
        got = e["eng"].uncertainty_from_logits(lg.to(DEV), mode).cpu().numpy()
        want = S.scores_from_prob(lg.softmax(2).mean(1), mode).numpy()
        np.testing.assert_allclose(got, want, rtol=2e-5)


def test_mc_dropout_matches_composed_oracle(env, golden):
    from oracle.synth import mc_masks, synth_pool_u8
    g = golden["mc_K3"]
    T, p, seed, start = int(g["T"]), float(g["p"]), int(g["seed"]), int(g["start"])
    eng = env[3]["eng"]
    imgs = synth_pool_u8(8, start=start).to(DEV)
    for mode in ("entropy", "margin", "confidence"):
        got = eng.scores(imgs, mode, T=T, drop_p=p, seed=seed, base_pos=0).cpu().numpy()
        want = g[f"scores/{mode}"]
        rel = np.abs(got - want) / np.abs(want)
        assert rel.max() < SCORE_RTOL, f"{mode}: {rel.max():.3e}"
    # device-generated masks == the oracle's explicit masks, bit for bit
    a = eng.scores(imgs, "entropy", T=T, drop_p=p, seed=seed, base_pos=5)
    b = eng.scores(imgs, "entropy", T=T, masks=mc_masks(seed, np.arange(5, 13), T, p))
    assert torch.equal(a, b)
    # keyed by pool position: scoring a sub-range reproduces the same numbers
    c = eng.scores(imgs[3:], "entropy", T=T, drop_p=p, seed=seed, base_pos=8)
    assert torch.equal(a[3:], c)


def test_hybrid_selector_matches_composed_oracle(env):
    import b200al
    from oracle import select_ref as S
    from oracle.synth import make_active
    e = env[2]
    L, k = e["L"], 8
    sel = b200al.B200HybridSelector(batch_size=8, num_workers=0, pin_memory=True, prefilter_factor=4,
                                    chunk_samples=32)
    ad = make_active(e["images"], L)
    got = sel.select_next_batch(ad, k, e["model"], torch.device(DEV))
    emb = e["g"]["embeddings"]
    want = S.hybrid_select(e["g"]["scores/entropy"], emb[:L], emb[L:], k, 4 * k)
    assert sorted(got) == sorted(_names(want, base=L))


def test_scores_are_chunk_invariant_and_deterministic(env):
    import b200al
    e = env[2]
    pool = e["images"][e["L"]:e["L"] + 24].to(DEV)
    a = e["eng"].scores(pool, "entropy")
    b = e["eng"].scores(pool, "entropy")
    assert torch.equal(a, b)
    small = b200al.Engine(DEV, 2, max_samples=5, precision=env["precision"])
    small.load_model(e["sd"])
    c = small.scores(pool, "entropy")
    assert torch.equal(a, c)
    # fp32 [n,1,H,W] input (what the reference's datasets hand over: to_tensor on the CPU) == uint8 input
    f32 = pool.cpu().float().div(255).unsqueeze(1).to(DEV)
    assert torch.equal(a, e["eng"].scores(f32, "entropy"))
    with pytest.raises(RuntimeError, match="max_samples"):
        from b200al import _lib
        import ctypes as C
        out = torch.empty(6, device=DEV)
        _lib.check(small.lib.b200al_unet_scores(small.ctx, C.c_void_p(pool.data_ptr()), 0, None, 6, 0, 1, 0.0, 0, 0, None,
                                                1e-8, C.c_void_p(out.data_ptr()), None, None), small.ctx)
    small.close()


def test_generic_dataloader_path_equals_resident_path(env):
    """A dataset without the resident uint8 fast path goes through a DataLoader
    like the reference (entropy_selector.py:30-35) and gives the same pick."""
    import b200al
    from oracle.synth import MirrorActive, MirrorExtendable, SynthBase
    e = env[3]

    class Plain(SynthBase):           # same samples, but no `images_u8` attribute to slice from
        def __init__(self, images):
            self._store = images
            self.names = SynthBase(images).names

        def get_sample(self, index, normalize=True):
            return SynthBase(self._store, self.names).get_sample(index, normalize)

    base = Plain(e["images"])
    names = list(base.names)
    ad = MirrorActive(MirrorExtendable(base, names[:e["L"]]), MirrorExtendable(base, names[e["L"]:]))
    sel = b200al.B200EntropySelector(batch_size=8, num_workers=0, pin_memory=False, chunk_samples=32)
    assert sel.select_next_batch(ad, e["k"], e["model"], torch.device(DEV)) == e["g"]["selected/entropy"].tolist()


def test_kcenter_fused_add_matches_oracle_and_reference_golden(golden):
    """CoresetSelector with loaded features, coreset_fusion="add" (coreset_selector.py:110-157): the two-matrix
    k-center kernels (global pair-distance sums + fused running min / sum) vs the streaming oracle on random
    matrices, and the drop-in selector end to end vs the reference's frozen selections (incl. the cold-start
    k-means++ branch over the loaded features)."""
    import os
    import b200al
    from oracle import select_ref as S
    from oracle.make_golden_fused import BUDGET, N_LAB, N_POOL, WEIGHT, loaded_features
    from oracle.synth import case_name, make_active, synth_pool_u8
    from oracle.unet_ref import OracleUNet, golden_state_dict
    rng = np.random.default_rng(21)
    eng = b200al.Engine(DEV, 2, max_samples=1)
    try:
        for n, d2, L, w in ((150, 200, 10, 0.3), (400, 96, 3, 0.8), (64, 1024, 20, 0.5)):
            feats = (rng.standard_normal((n, 512)) * 0.3 + rng.standard_normal((1, 512))).astype(np.float32)
            loaded = (rng.standard_normal((n, d2)) + rng.standard_normal((6, d2))[rng.integers(0, 6, n)]).astype(np.float32)
            for metric in ("l2", "cosine"):
                for crit in ("min", "mean"):
                    want, gaps = S.kcenter_fused_add_stream(feats, loaded, w, L, 12, metric, crit, return_gaps=True)
                    got = eng.kcenter_fused(torch.from_numpy(feats), torch.from_numpy(loaded), w, L, 12, metric, crit)
                    got = got.cpu().tolist()
                    # (cosine dot products come from an fp32 BLAS GEMM in the oracle: a margin is required, and all
                    # twelve runs have it - smallest top-2 gap 3.0e-5 - so every one is compared pick by pick)
                    assert min(gaps) > 1e-5, (n, d2, metric, crit, min(gaps))
                    assert got == want, (n, d2, metric, crit, min(gaps))
                    assert len(set(got)) == 12 and min(got) >= L
    finally:
        eng.close()
    gf = np.load(os.path.join(os.path.dirname(__file__), "golden", "coreset_fused.npz"))
    names = [case_name(i) for i in range(N_LAB + N_POOL)]
    fd = loaded_features(names)
    model = OracleUNet(golden_state_dict(3, golden["model"]))
    images = synth_pool_u8(N_LAB + N_POOL)
    kw = dict(batch_size=8, num_workers=0, pin_memory=False, feature_dict=fd, loaded_feature_weight=WEIGHT)
    for metric in ("l2", "cosine"):
        for crit in ("min", "mean"):
            sel = b200al.B200CoresetSelector(metric=metric, coreset_criteria=crit, coreset_fusion="add", **kw)
            got = sel.select_next_batch(make_active(images, N_LAB), BUDGET, model, torch.device(DEV))
            assert sorted(got) == gf[f"selected/{metric}/{crit}"].tolist(), (metric, crit)
    np.random.seed(int(gf["seed/cold"]))
    cold = b200al.B200CoresetSelector(metric="l2", **kw).select_next_batch(make_active(images, 0), BUDGET, model,
                                                                          torch.device(DEV))
    assert cold == gf["selected/cold"].tolist()
    with pytest.raises(NotImplementedError):   # "cat" is built since round 2 (tests/test_gpu_round2.py); anything else is not
        b200al.B200CoresetSelector(batch_size=1, num_workers=0, feature_dict=fd, coreset_fusion="mul")
