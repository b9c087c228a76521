"""Pins the oracle against the UNMODIFIED reference run here (container only;
skipped on the GPU box where /root/reference does not exist)."""
import numpy as np
import pytest
import torch

from oracle.ref_import import reference_available

pytestmark = [pytest.mark.reference,
              pytest.mark.skipif(not reference_available(), reason="/root/reference not present")]


@pytest.fixture(scope="module")
def ref():
    from oracle.ref_import import load_reference
    return load_reference()


def test_seeded_weights_identical_to_reference_unet(ref):
    from oracle.ref_import import build_reference_unet
    from oracle.unet_ref import make_state_dict
    torch.manual_seed(1337)
    rsd = build_reference_unet(3).state_dict()
    sd = make_state_dict(1337, 3)
    assert list(rsd.keys()) == list(sd.keys())
    assert all(torch.equal(rsd[k], sd[k]) for k in sd)


def test_forward_and_embedding_identical(ref, golden):
    from oracle.ref_import import build_reference_unet
    from oracle.synth import synth_pool_u8
    from oracle.unet_ref import golden_state_dict, unet_embed, unet_forward
    sd = golden_state_dict(3, golden["model"])
    m = build_reference_unet(3)
    m.load_state_dict(sd)
    m.eval()
    x = synth_pool_u8(2, start=77).float().div(255).unsqueeze(1)
    with torch.no_grad():
        assert torch.equal(m(x), unet_forward(sd, x))
        assert torch.equal(m.get_enc_feature(x), unet_embed(sd, x))


def test_mc_masks_equal_dropout_modules_replaced(ref, golden):
    """Composed MC oracle == reference module with each Dropout2d swapped for
    an explicit-mask multiply."""
    from oracle.ref_import import build_reference_unet
    from oracle.synth import mc_masks, synth_pool_u8
    from oracle.unet_ref import dropout_sites, golden_state_dict, unet_forward
    sd = golden_state_dict(3, golden["model"])
    m = build_reference_unet(3, dropout_prob=0.1)
    m.load_state_dict(sd)
    m.eval()
    masks = mc_masks(11, [3], 1, 0.1)[:, 0]
    sites, total = dropout_sites()
    assert total == 2944

    class Mul(torch.nn.Module):
        def __init__(self, v):
            super().__init__()
            self.v = v

        def forward(self, x):
            return x * self.v[:, :, None, None]

    for prefix, off, c in sites:
        blk = m.get_submodule(prefix)
        assert isinstance(blk[1], torch.nn.Dropout2d)
        blk[1] = Mul(masks[:, off:off + c])
    x = synth_pool_u8(1, start=3).float().div(255).unsqueeze(1)
    with torch.no_grad():
        np.testing.assert_allclose(m(x).numpy(), unet_forward(sd, x, masks=masks).numpy(), rtol=0, atol=1e-6)


def test_kcenter_restatements_equal_reference(ref):
    from sklearn.metrics import pairwise_distances
    from oracle import select_ref as S
    rng = np.random.default_rng(3)
    for metric in ("l2", "cosine"):
        feats = (rng.standard_normal((500, 512)) * 0.3 + 0.2).astype(np.float32)
        dist = pairwise_distances(feats, metric=metric)
        want = ref.kcenter_greedy(dist, 500, 40, np.arange(25), "min")
        assert sorted(S.kcenter_greedy_matrix(dist, 40, np.arange(25))) == sorted(int(i) for i in want)
        assert sorted(S.kcenter_greedy_stream(feats, 25, 40, metric)) == sorted(int(i) for i in want)
        # the reference's /sum normalisation (coreset_selector.py:153) does not change the picks
        d2 = dist / dist.sum()
        assert sorted(int(i) for i in ref.kcenter_greedy(d2, 500, 40, np.arange(25), "min")) == sorted(int(i) for i in want)
        # criteria "mean" (coreset_selector.py:35-36): matrix restatement and the running-sum form
        want_m = sorted(int(i) for i in ref.kcenter_greedy(dist, 500, 40, np.arange(25), "mean"))
        assert sorted(S.kcenter_greedy_matrix(dist, 40, np.arange(25), "mean")) == want_m
        got_m, gaps = S.kcenter_greedy_mean_stream(feats, 25, 40, metric, return_gaps=True)
        if min(gaps) > 1e-5 * float(dist.mean()):   # no fp32-rounding-level tie between the two best rows
            assert sorted(got_m) == want_m
    col = S.l2_rows(feats, feats[7], 7)
    np.testing.assert_array_equal(col, pairwise_distances(feats, metric="l2")[:, 7])


def test_reference_selector_on_mirror_datasets(ref, golden):
    """The reference's EntropySelector accepts the mirror dataset types the GPU
    tests use, and returns the golden selection."""
    from oracle.ref_import import build_reference_unet
    from oracle.synth import make_active, synth_pool_u8
    from oracle.unet_ref import golden_state_dict
    g = golden["pool_K3"]
    L, n = int(g["n_labeled"]), 6
    sd = golden_state_dict(3, golden["model"])
    m = build_reference_unet(3)
    m.load_state_dict(sd)
    ad = make_active(synth_pool_u8(L + n), L)
    sel = ref.EntropySelector(batch_size=4, num_workers=0, pin_memory=False)
    scores, names = sel.cal_scores(ad, m, torch.device("cpu"))
    np.testing.assert_allclose(torch.stack(scores).numpy(), g["scores/entropy"][:n], rtol=1e-5)
    assert names == ad.get_pool_dataset().image_idx


def test_antialiased_resize_matches_torch_and_torchvision(ref):
    """The third-party algorithm behind JointResize (joint_transform.py:24): torch interpolate(antialias=True)."""
    import torch.nn.functional as TF
    import torchvision.transforms.functional as VF
    from oracle.prep_ref import resize_bilinear_aa
    rng = np.random.default_rng(3)
    for hs, ws in [(336, 544), (512, 512), (256, 256), (100, 90), (257, 255), (1900, 2600), (256, 700), (31, 1024)]:
        x = rng.random((hs, ws), dtype=np.float32)
        want = TF.interpolate(torch.from_numpy(x)[None, None], size=(256, 256), mode="bilinear", align_corners=False,
                              antialias=True)[0, 0].numpy()
        tv = VF.resize(torch.from_numpy(x)[None], [256, 256], VF.InterpolationMode.BILINEAR)[0].numpy()
        got = resize_bilinear_aa(x, (256, 256))
        assert np.abs(got - want).max() < 1e-6, (hs, ws)
        assert np.abs(got - tv).max() < 1e-6, (hs, ws)


@pytest.mark.parametrize("normalize", [False, True])
def test_input_pipeline_matches_reference_fugc_dataset(ref, tmp_path, normalize):
    """PNG files -> the reference's FUGCDataset exactly as al_train builds the pool dataset -> oracle restatement;
    and the pool cache's raw reader sees the same pixels the reference decodes."""
    from b200al.pool_cache import DevicePoolCache, _raw_reader
    from oracle.make_golden_prep import write_fugc_tree
    from oracle.prep_ref import prep_image, synth_raw_u8
    raws = [synth_raw_u8(i) for i in range(3, 8)]
    names = [f"c{i}" for i in range(len(raws))]
    write_fugc_tree(str(tmp_path), raws, names)
    ds = ref.FUGCDataset(data_path=str(tmp_path), split="train", normalize=ref.ZScoreNormalize() if normalize else None,
                         transform=None, image_channels=1, image_size=256)
    read = _raw_reader(ds)
    assert read is not None
    for i in range(len(ds)):
        smp = ds.get_sample(i)
        k = names.index(smp["case_name"])
        assert np.array_equal(read(i), raws[k])
        got = prep_image(raws[k], (256, 256), normalize)
        assert smp["image"].shape == (1, 256, 256)
        assert np.abs(got - smp["image"][0].numpy()).max() < (5e-6 if normalize else 1e-6)
    cache = DevicePoolCache.try_build(ds, "cpu")  # host-side bookkeeping only (no kernels on a CPU device)
    assert cache is not None and cache.n == len(ds) and cache.normalize == normalize and cache.out_hw == (256, 256)
    ds.transform = object()  # a dataset with (random) transforms must never be cached
    assert DevicePoolCache.try_build(ds, "cpu") is None


def test_hook_against_the_real_al_trainer(ref, monkeypatch):
    """`install_into_al_trainer` on the reference's own ALTrainer / ALConfig (al_trainer.py:87,291,802-883): b200-*
    names build the drop-in selectors from the real config object, the reference's names and its ValueError for
    unknown names are untouched, and the selectors pass isinstance against the reference ABC."""
    import importlib
    import sys
    import b200al
    from oracle.ref_import import REFERENCE_SRC
    sys.path.insert(0, REFERENCE_SRC)
    try:
        m = importlib.import_module("training.al_trainer")
    finally:
        sys.path.remove(REFERENCE_SRC)
    monkeypatch.setattr(m.ALTrainer, "_setup_active_selector", m.ALTrainer._setup_active_selector)  # restored after
    b200al.install_into_al_trainer(m.ALTrainer, reference_abc=ref.ActiveSelector)
    cfg = m.ALConfig(active_selector_name="b200-coreset-l2", coreset_criteria="mean", batch_size=7, num_workers=0,
                     pin_memory=False)
    t = object.__new__(m.ALTrainer)  # no work dir / logger / model needed for selector set-up
    t.config = cfg
    t._setup_loss()  # the real DiceAndCELoss (al_trainer.py:782-798): b200-badge reads its dice / ce modules
    t._setup_active_selector()
    sel = t.active_selector
    assert isinstance(sel, b200al.B200CoresetSelector) and isinstance(sel, ref.ActiveSelector)
    assert (sel.batch_size, sel.num_workers, sel.pin_memory, sel.metric, sel.coreset_criteria) == (7, 0, False, "l2", "mean")
    for name in b200al.SELECTORS:
        cfg.active_selector_name = name
        t._setup_active_selector()
        assert isinstance(t.active_selector, ref.ActiveSelector)
    cfg.active_selector_name = "b200-badge"
    t._setup_active_selector()
    assert (t.active_selector.dice_smooth, t.active_selector.dice_do_bg) == (1e-5, True)
    cfg.active_selector_name = "entropy"
    t._setup_active_selector()
    assert type(t.active_selector) is ref.EntropySelector
    cfg.active_selector_name = "no-such-selector"
    with pytest.raises(ValueError):
        t._setup_active_selector()


def test_kmean_oracle_matches_reference_selector_live(ref, golden):
    """Live run of the reference KMeanSelector (not the frozen golden): fresh seed, K=3 pool."""
    from oracle import select_ref as S
    from oracle.kmean_ref import kmean_select
    from oracle.ref_import import build_reference_unet
    from oracle.synth import case_name, make_active, synth_pool_u8
    from oracle.unet_ref import golden_state_dict
    sd = golden_state_dict(3, golden["model"])
    model = build_reference_unet(3)
    model.load_state_dict(sd)
    images = synth_pool_u8(10 + 40, start=300)
    emb = S.pool_embeddings(sd, images)
    for kw in (dict(metric="cosine"), dict(metric="l2", coreset_criteria="mean", sharp_factor=2.0)):
        sel = ref.KMeanSelector(batch_size=8, num_workers=0, pin_memory=False, **kw)
        np.random.seed(99)
        want = sel.select_next_batch(make_active(images, 10, reference=ref), 7, model, torch.device("cpu"))
        idx = kmean_select(emb[10:], emb[:10], 7, kw["metric"], kw.get("coreset_criteria", "min"),
                           kw.get("sharp_factor", 1.0), False, np.random.RandomState(99))
        names = make_active(images, 10).get_pool_dataset().image_idx
        assert sorted(want) == sorted({names[int(i)] for i in idx})


def test_badge_oracle_bit_identical_to_reference_cal_scores(ref, golden):
    import importlib
    import sys
    from oracle.badge_ref import grad_embedding_autograd
    from oracle.ref_import import REFERENCE_SRC, build_reference_unet
    from oracle.synth import make_active, synth_pool_u8
    from oracle.unet_ref import golden_state_dict
    sys.path.insert(0, REFERENCE_SRC)
    try:
        al = importlib.import_module("activelearning")
        dl = importlib.import_module("losses.dice_loss")
    finally:
        sys.path.remove(REFERENCE_SRC)
    sd = golden_state_dict(3, golden["model"])
    model = build_reference_unet(3)
    model.load_state_dict(sd)
    images = synth_pool_u8(2 + 3, start=700)
    for do_bg in (True, False):
        dice = dl.DiceLoss(num_classes=2, smooth=1e-5, do_bg=do_bg, softmax=True, batch=False, squared=False)
        sel = al.BADGESelector(dice_loss=dice, ce_loss=torch.nn.CrossEntropyLoss(), batch_size=1, num_workers=0,
                               pin_memory=False)
        _names, emb = sel.cal_scores(make_active(images, 2, reference=ref), model, torch.device("cpu"))
        assert np.array_equal(emb, grad_embedding_autograd(sd, images[2:].float().div(255).unsqueeze(1), 1e-5, do_bg))


def test_fused_coreset_oracle_matches_reference_live(ref, golden, tmp_path, monkeypatch):
    """Unmodified reference CoresetSelector with loaded features (coreset_fusion="add") vs the streaming oracle, on a
    pool / weight / seed different from the frozen golden.  h5py is absent from this image: the reference module's
    `h5py` attribute is pointed at oracle.make_golden_fused.H5Shim (File + Dataset over numpy files) for the read."""
    import importlib
    from oracle import select_ref as S
    from oracle.make_golden_fused import H5Shim, write_feature_files
    from oracle.ref_import import build_reference_unet
    from oracle.synth import make_active, synth_pool_u8
    from oracle.unet_ref import golden_state_dict
    cs = importlib.import_module("activelearning.coreset_selector")
    monkeypatch.setattr(cs, "h5py", H5Shim)
    sd = golden_state_dict(3, golden["model"])
    model = build_reference_unet(3)
    model.load_state_dict(sd)
    images = synth_pool_u8(5 + 25, start=400)
    ad = make_active(images, 5, reference=ref)
    names = list(ad.get_train_dataset().image_idx) + list(ad.get_pool_dataset().image_idx)
    rng = np.random.default_rng(17)
    feats = {n: rng.standard_normal(96).astype(np.float32) + (i % 3) for i, n in enumerate(names)}
    write_feature_files(str(tmp_path), feats)
    loaded = np.stack([feats[n] for n in names])
    emb = S.pool_embeddings(sd, images)
    for metric, crit, w in (("l2", "min", 0.7), ("cosine", "mean", 0.25)):
        sel = ref.CoresetSelector(batch_size=8, num_workers=0, pin_memory=False, metric=metric, coreset_criteria=crit,
                                  coreset_fusion="add", feature_path=str(tmp_path), loaded_feature_weight=w)
        want = sel.select_next_batch(make_active(images, 5, reference=ref), 6, model, torch.device("cpu"))
        picks = S.kcenter_fused_add_stream(emb, loaded, w, 5, 6, metric, crit)
        assert sorted(map(str, want)) == sorted(names[i] for i in picks)


def test_dice_oracle_matches_reference_diceloss_class(ref):
    """Dice-uncertainty = the reference's DiceLoss (src/losses/dice_loss.py:32-72, configured like
    al_trainer.py:785-791 but do_bg=True) of each image's softmax map against its own argmax."""
    from oracle import select_ref as S
    g = torch.Generator().manual_seed(5)
    for K in (2, 3, 4):
        logits = torch.randn(5, K, 40, 24, generator=g) * 2
        loss = ref.DiceLoss(num_classes=K - 1, smooth=1e-5, do_bg=True, softmax=True)
        want = torch.stack([loss(logits[i:i + 1], logits[i:i + 1].softmax(1).argmax(1)) for i in range(5)])
        got = S.dice_scores(logits)
        np.testing.assert_allclose(got.numpy(), want.numpy(), rtol=1e-6, atol=1e-7)


def test_cat_fusion_restatement_matches_reference_live(ref, golden, tmp_path, monkeypatch):
    """coreset_fusion="cat" run live (numpy 2 here -> float64 product) == the restatement under the same rule; the
    product path implements the reference's pinned numpy < 2 rule (float32), which cannot be run live here."""
    import importlib
    from oracle import select_ref as S
    from oracle.make_golden_fused import H5Shim, write_feature_files
    from oracle.ref_import import build_reference_unet
    from oracle.synth import make_active, synth_pool_u8
    from oracle.unet_ref import golden_state_dict
    cs = importlib.import_module("activelearning.coreset_selector")
    monkeypatch.setattr(cs, "h5py", H5Shim, raising=False)
    sd = golden_state_dict(3, golden["model"])
    model = build_reference_unet(3)
    model.load_state_dict(sd)
    images = synth_pool_u8(4 + 14, start=300)
    ad = make_active(images, 4, reference=ref)
    names = list(ad.get_train_dataset().image_idx) + list(ad.get_pool_dataset().image_idx)
    rng = np.random.default_rng(3)
    fd = {n: rng.standard_normal(96).astype(np.float32) for n in names}
    write_feature_files(str(tmp_path), fd)
    loaded = np.stack([fd[n] for n in names])
    emb = S.pool_embeddings(sd, images)
    for metric in ("l2", "cosine"):
        sel = ref.CoresetSelector(batch_size=8, num_workers=0, pin_memory=False, metric=metric, coreset_fusion="cat",
                                  feature_path=str(tmp_path), loaded_feature_weight=0.6)
        live = sel.select_next_batch(make_active(images, 4, reference=ref), 5, model, torch.device("cpu"))
        rule = "nep50" if int(np.__version__.split(".")[0]) >= 2 else "legacy"
        picks = S.coreset_cat_select(emb, loaded, 0.6, 4, 5, metric, "min", rule)
        assert sorted(map(str, live)) == sorted(names[i] for i in picks)
