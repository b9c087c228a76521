"""Full-size and cross-path checks on the GPU that do not need a CPU oracle run:
size-independent properties at BASELINE.json's sizes (100k-row k-center, 10k-image
pools), equivalence of the specialised kernels (column-folded conv, fused head)
with the generic tcgen05 path, and the selectors' edge cases."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _engine(K, env=None, **kw):
    """Engine built under temporary B200AL_* planning switches (read at ctx creation)."""
    import b200al
    saved = {}
    for k, v in (env or {}).items():
        saved[k] = os.environ.get(k)
        os.environ[k] = v
    try:
        return b200al.Engine(DEV, K, **kw)
    finally:
        for k, v in saved.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


@pytest.mark.parametrize("K", [2, 3])
def test_specialised_kernels_equal_generic_path(golden, K):
    """Folded conv + fused head + super-tiles + CTA pairs vs. the plain generic per-tap kernel + separate head: same
    operands, different summation order only -> scores agree to 1e-5 relative, logits to 1e-3 absolute."""
    from oracle.synth import synth_pool_u8
    from oracle.unet_ref import golden_state_dict
    sd = golden_state_dict(K, golden["model"])
    imgs = synth_pool_u8(12, start=500).to(DEV)
    fast = _engine(K, max_samples=8)
    slow = _engine(K, {"B200AL_FOLD": "0", "B200AL_FUSE_HEAD": "0", "B200AL_MT2": "0", "B200AL_PAIR": "0"}, max_samples=8)
    try:
        fast.load_model(sd)
        slow.load_model(sd)
        for mode in ("entropy", "margin", "confidence"):
            a, la = fast.scores(imgs, mode, return_logits=True)
            b, lb = slow.scores(imgs, mode, return_logits=True)
            rel = ((a - b).abs() / b.abs()).max().item()
            assert rel < 1e-5, (mode, rel)
            # the unfused head reads the last feature map back as a bf16 hi/lo pair (16 bits), the fused one keeps fp32
            assert (la - lb).abs().max().item() < 1e-3
        # MC-dropout masks reach the folded layers through the same per-sample table
        a = fast.scores(imgs, "entropy", T=4, drop_p=0.1, seed=3)
        b = slow.scores(imgs, "entropy", T=4, drop_p=0.1, seed=3)
        assert ((a - b).abs() / b.abs()).max().item() < 1e-5
        assert (fast.embed(imgs) - slow.embed(imgs)).abs().max().item() < 1e-4
    finally:
        fast.close()
        slow.close()


def test_kcenter_100k_rows_greedy_invariants():
    """Config 3 size: 100k x 512 embeddings, 500 initial centres.  Every pick must be the arg-max of the
    running min-distance at its step (checked against torch fp64 distances for the recorded values),
    picks are distinct pool rows, and the recorded farthest distances are non-increasing."""
    import b200al
    eng = b200al.Engine(DEV, 2, max_samples=1)
    g = torch.Generator(device=DEV).manual_seed(11)
    n, L, budget = 100_000, 500, 64
    feats = torch.randn((n, 512), device=DEV, generator=g) * 0.25
    picks, vals = eng.kcenter(feats, L, budget, "l2", return_vals=True)
    p = picks.cpu().tolist()
    assert len(set(p)) == budget and min(p) >= L and max(p) < n
    v = vals.cpu().numpy()
    assert np.all(np.diff(v) <= 1e-6), "farthest-point distances must not increase"
    # replay: min distance of every row to the current centre set, in fp64 on the GPU
    f64 = feats.double()
    mind = torch.cdist(f64, f64[:L]).min(dim=1).values
    mind[:L] = -1
    for step, (row, val) in enumerate(zip(p, v)):
        best = mind.max().item()
        assert abs(mind[row].item() - best) <= 1e-5 * best, f"step {step}: row {row} is not a farthest point"
        assert abs(val - best) <= 1e-5 * best
        mind = torch.minimum(mind, (f64 - f64[row]).norm(dim=1))
        mind[row] = -1
    eng.close()


def test_topk_one_million_scores_with_duplicates():
    import b200al
    eng = b200al.Engine(DEV, 2, max_samples=1)
    g = torch.Generator(device=DEV).manual_seed(5)
    sc = torch.randint(0, 50_000, (1_000_000,), device=DEV, generator=g).float() / 7.0
    vals, pos = eng.topk(sc, 500, base_pos=0)
    order = torch.sort(sc.cpu(), descending=True, stable=True).indices[:500]
    assert pos.cpu().tolist() == order.tolist()
    assert torch.equal(vals.cpu(), sc.cpu()[order])
    eng.close()


def test_pool_10k_mc_scores_shard_and_chunk_invariant(golden):
    """Config 2 shape at reduced length (the property is per image): scores of a pool slice do not depend
    on how the pool is chunked or sharded, because masks are keyed by pool position."""
    from oracle.synth import synth_pool_u8
    from oracle.unet_ref import golden_state_dict
    sd = golden_state_dict(3, golden["model"])
    imgs = synth_pool_u8(48, start=2000).to(DEV)
    a_eng = _engine(3, max_samples=128)
    b_eng = _engine(3, max_samples=24)
    try:
        a_eng.load_model(sd)
        b_eng.load_model(sd)
        whole = a_eng.scores(imgs, "entropy", T=8, drop_p=0.1, seed=2025, base_pos=7000)
        parts = torch.cat([b_eng.scores(imgs[i:i + 16], "entropy", T=8, drop_p=0.1, seed=2025, base_pos=7000 + i)
                           for i in range(0, 48, 16)])
        assert torch.equal(whole, parts)
    finally:
        a_eng.close()
        b_eng.close()


def test_selector_edge_cases(golden):
    import b200al
    from oracle.synth import make_active, synth_pool_u8
    from oracle.unet_ref import OracleUNet, golden_state_dict
    model = OracleUNet(golden_state_dict(2, golden["model"]))
    images = synth_pool_u8(9)
    ad = make_active(images, 4)  # 4 labeled, 5 in the pool
    pool_names = list(ad.get_pool_dataset().image_idx)
    for cls in (b200al.B200EntropySelector, b200al.B200MarginSelector, b200al.B200ConfidenceSelector):
        sel = cls(batch_size=3, num_workers=0, pin_memory=False)
        assert sel.select_next_batch(ad, 0, model, torch.device(DEV)) == []
        got = sel.select_next_batch(ad, 50, model, torch.device(DEV))  # budget larger than the pool: the whole pool
        assert sorted(got) == sorted(pool_names) and len(set(got)) == len(got)
    cs = b200al.B200CoresetSelector(batch_size=3, num_workers=0, pin_memory=False, metric="l2")
    got = cs.select_next_batch(ad, 50, model, torch.device(DEV))
    assert sorted(got) == sorted(pool_names)
    # cold start (no labeled data): the reference's seeded random branch
    ad0 = make_active(images, 0)
    torch.manual_seed(3)
    a = b200al.B200EntropySelector(batch_size=3, num_workers=0).select_next_batch(ad0, 4, model, torch.device(DEV))
    torch.manual_seed(3)
    _, idx = torch.sort(torch.rand(9), descending=True)
    assert a == [ad0.get_pool_dataset().image_idx[i] for i in idx[:4]]
    with pytest.raises(RuntimeError):
        b200al.B200CoresetSelector(batch_size=1, num_workers=0, coreset_criteria="median")
