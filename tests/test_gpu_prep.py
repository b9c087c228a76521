"""Input pipeline on the GPU: b200al_prep_images (to_tensor -> JointResize -> ZScoreNormalize
statistics) and the first conv's fused normalisation vs the oracle restatement (oracle/prep_ref.py, pinned against
the reference's FUGCDataset / torchvision) and the frozen reference outputs in tests/golden/prep.npz; the selectors
through the device pool cache vs the same selectors fed by the CPU pipeline through a DataLoader."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _pack(raws, dev=DEV):
    sizes = np.array([r.shape for r in raws], np.int64)
    nb = sizes[:, 0] * sizes[:, 1]
    offs = np.concatenate([[0], np.cumsum(nb)[:-1]]).astype(np.int64)
    raw = torch.from_numpy(np.concatenate([r.reshape(-1) for r in raws])).to(dev)
    return raw, torch.from_numpy(offs).to(dev), torch.from_numpy(sizes.astype(np.int32)).to(dev), \
        int(sizes[:, 0].max()), int(sizes[:, 1].max())


def test_prep_matches_oracle_and_reference_golden():
    import b200al
    from oracle.prep_ref import prep_image, synth_raw_u8
    rng = np.random.default_rng(9)
    raws = [synth_raw_u8(i) for i in range(10)]
    raws.append(rng.integers(0, 256, (256, 256), dtype=np.uint8))     # already model-sized: identity
    raws.append(rng.integers(0, 256, (1500, 2100), dtype=np.uint8))   # 8x down-scale, 19 taps per axis
    raws.append(rng.integers(0, 256, (40, 33), dtype=np.uint8))       # strong up-sampling
    raws.append(np.full((300, 300), 77, np.uint8))                    # constant image: std clipped at 1e-8
    eng = b200al.Engine(DEV, 2, max_samples=8)  # 14 images > max_samples: exercises the internal chunking
    try:
        raw, offs, hw, mh, mw = _pack(raws)
        imgs, norm = eng.prep(raw, offs, hw, mh, mw, normalize=True)
        imgs2, none = eng.prep(raw, offs, hw, mh, mw, normalize=False)
        assert none is None and torch.equal(imgs, imgs2)
        imgs, norm = imgs.cpu().numpy(), norm.cpu().numpy()
        for i, r in enumerate(raws):
            want = prep_image(r, (256, 256), False)
            assert np.abs(imgs[i] - want).max() < 1e-6, i
            m, s = want.astype(np.float64).mean(), want.astype(np.float64).std(ddof=1)
            assert abs(norm[i, 0] - m) <= 1e-6 * max(abs(m), 1e-3), i
            if i != 13:  # the constant image's std is rounding noise of the filter weights on both sides
                assert abs(norm[i, 1] - s) <= 1e-5 * s, i
        assert np.array_equal(imgs[10], raws[10].astype(np.float32) / np.float32(255))  # identity resize is exact
        assert norm[13, 1] <= 1e-6  # constant image
        # frozen outputs of the reference dataset pipeline
        g = np.load(os.path.join(os.path.dirname(__file__), "golden", "prep.npz"))
        for i in range(3):
            assert np.abs(imgs[i] - g["resized"][i]).max() < 1e-6
            assert abs(norm[i, 0] - g["mean"][i]) < 1e-6 and abs(norm[i, 1] - g["std"][i]) < 1e-6
            z = (imgs[i] - norm[i, 0]) / norm[i, 1]
            assert np.abs(z[::16, ::16] - g["zprobe"][i]).max() < 1e-5
        with pytest.raises(RuntimeError):  # more than 17x down-scale is rejected, not silently truncated
            eng.prep(raw, offs, hw, 256 * 18, mw, normalize=False)
    finally:
        eng.close()


@pytest.mark.parametrize("K", [2, 3])
def test_fused_zscore_equals_explicitly_normalised_input(golden, K):
    """(x - mean) / std inside the first conv's loader == feeding the normalised fp32 image."""
    import b200al
    from oracle.prep_ref import synth_raw_u8
    from oracle.unet_ref import golden_state_dict
    raws = [synth_raw_u8(i) for i in range(20, 26)]
    eng = b200al.Engine(DEV, K, max_samples=8)
    try:
        eng.load_model(golden_state_dict(K, golden["model"]))
        imgs, norm = eng.prep(*_pack(raws), normalize=True)
        z = (imgs - norm[:, 0, None, None]) / norm[:, 1, None, None]
        for mode in ("entropy", "margin", "confidence"):
            a, la = eng.scores(imgs, mode, norm=norm, return_logits=True)
            b, lb = eng.scores(z, mode, return_logits=True)
            assert (la - lb).abs().max().item() < 1e-4
            assert ((a - b).abs() / b.abs()).max().item() < 1e-4
        assert (eng.embed(imgs, norm=norm) - eng.embed(z)).abs().max().item() < 1e-4
        # and it differs from the un-normalised scores (the normalisation is really applied)
        assert (eng.scores(imgs, "entropy", norm=norm) - eng.scores(imgs, "entropy")).abs().max().item() > 1e-4
    finally:
        eng.close()


@pytest.mark.parametrize("normalize", [False, True])
def test_selectors_through_pool_cache_equal_dataloader_path_and_oracle(golden, normalize):
    """--do-normalize on/off.  Pool cache (GPU resize + fused z-score) vs pool_cache='off' (CPU pipeline through the
    reference's DataLoader settings) vs the CPU oracle end to end: same selected names, scores within 1e-3."""
    import b200al
    from oracle import select_ref as S
    from oracle.prep_ref import synth_raw_u8
    from oracle.synth import MirrorActive, MirrorExtendable, SynthRawBase
    from oracle.unet_ref import OracleUNet, golden_state_dict, unet_forward
    sd = golden_state_dict(3, golden["model"])
    model = OracleUNet(sd)
    raws = [synth_raw_u8(i) for i in range(100, 118)]

    def active():
        base = SynthRawBase(raws, 256, normalize)
        return MirrorActive(MirrorExtendable(base, base.names[:4]), MirrorExtendable(base, base.names[4:]))

    dev = torch.device(DEV)
    kw = dict(batch_size=4, num_workers=0, pin_memory=False)
    ad = active()
    pool = ad.get_pool_dataset()
    x = torch.stack([pool.get_sample(i)["image"] for i in range(len(pool))])  # oracle pipeline on the CPU
    with torch.no_grad():
        logits = unet_forward(sd, x)
    for name, cls in (("entropy", b200al.B200EntropySelector), ("margin", b200al.B200MarginSelector),
                      ("confidence", b200al.B200ConfidenceSelector)):
        want = S.SCORE_FNS[name](logits).numpy()
        cached, loader = cls(**kw), cls(pool_cache="off", **kw)
        sc, names = cached.cal_scores(active(), model, dev)
        sl, _ = loader.cal_scores(active(), model, dev)
        assert list(names) == pool.image_idx
        assert np.abs(sc.cpu().numpy() - want).max() <= 1e-3 * np.abs(want).max()
        assert np.abs(sl.cpu().numpy() - want).max() <= 1e-3 * np.abs(want).max()
        a = cached.select_next_batch(active(), 5, model, dev)
        b = loader.select_next_batch(active(), 5, model, dev)
        assert a == b == [pool.image_idx[i] for i in S.rank_desc_stable(want, 5)]
        assert cached.last_timing["pool"] == len(pool) and cached.last_timing["selected"] == 5
    # coreset: pool rows come from the cache, labeled rows too (no transforms here) - same picks either way
    a = b200al.B200CoresetSelector(metric="l2", **kw).select_next_batch(active(), 5, model, dev)
    b = b200al.B200CoresetSelector(metric="l2", pool_cache="off", **kw).select_next_batch(active(), 5, model, dev)
    assert a == b and len(set(a)) == 5
    # the cache survives the pool shrinking between rounds (extend_train_set only edits the name lists)
    ad = active()
    sel = b200al.B200EntropySelector(**kw)
    first = sel.select_next_batch(ad, 3, model, dev)
    cache = ad.get_pool_dataset().dataset._b200_pool_cache
    ad.extend_train_set(first)
    second = sel.select_next_batch(ad, 3, model, dev)
    assert ad.get_pool_dataset().dataset._b200_pool_cache is cache
    assert not set(first) & set(second)
