"""precision="mixed": the deep layers of the SCORING path contract single fp16 operands (conv_tc.cuh FMT 1/2/3), the rest
stays bf16x3.  Same parity bars as the bf16x3 path (logits 2e-2 abs, scores 1e-3 rel, identical selection), plus a
comparison with the CPU emulation of exactly this mix (tools/precision_table.py) that pins the arithmetic."""
import os
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

DEV = "cuda:0"
LOGIT_ATOL, SCORE_RTOL = 2e-2, 1e-3
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module", params=["mixed", "mixed-deep"])
def mixed(request, golden):
    import b200al
    from oracle.unet_ref import golden_state_dict
    e = {"precision": request.param}
    for K in (2, 3):
        eng = b200al.Engine(DEV, K, max_samples=32, precision=request.param)
        sd = golden_state_dict(K, golden["model"])
        eng.load_model(sd)
        e[K] = (eng, sd)
    yield e
    for K in (2, 3):
        e[K][0].close()


@pytest.mark.parametrize("K", [2, 3])
def test_mixed_logits_and_scores_match_reference(mixed, golden, K):
    from oracle.synth import synth_pool_u8
    eng, _sd = mixed[K]
    g = golden[f"pool_K{K}"]
    L, n = int(g["n_labeled"]), int(g["n_pool"])
    pool = synth_pool_u8(L + n)[L:].to(DEV)
    _s, logits = eng.scores(pool[:2], "entropy", return_logits=True)
    err = np.abs(logits.cpu().numpy() - g["probe_logits"]).max()
    print(f"{mixed['precision']} K={K}: max |logit err| {err:.3e}")
    assert err < LOGIT_ATOL, f"max abs logit err {err:.3e}"
    assert err < 0.6 * LOGIT_ATOL, "less than 40 % head-room under the logit tolerance"
    for mode in ("entropy", "margin", "confidence"):
        got = eng.scores(pool, mode).cpu().numpy()
        want = g[f"scores/{mode}"]
        rel = (np.abs(got - want) / np.abs(want)).max()
        assert rel < 0.1 * SCORE_RTOL, f"{mode}: max rel score err {rel:.3e}"


def test_mixed_matches_cpu_emulation_of_the_same_mix(mixed):
    """The kernels do what tools/precision_table.py emulates: fp16-rounded operands in the group, fp16 storage inside
    it, bf16 hi/lo elsewhere, fp32 accumulation."""
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import precision_table as P
    from oracle.synth import synth_pool_u8
    eng, sd = mixed[3]
    imgs = synth_pool_u8(2, start=16)
    fm, sto = P.mix_formats("wide" if mixed["precision"] == "mixed" else "deep")
    with torch.no_grad():
        want = P.forward(sd, imgs.float().div(255).unsqueeze(1), fm, sto)
    _s, got = eng.scores(imgs.to(DEV), "entropy", return_logits=True)
    err = (got.cpu() - want).abs().max().item()
    # not bit-level: the GPU and the CPU sum in different fp32 orders, and wherever that moves an activation across an
    # fp16 rounding boundary the two diverge by an fp16 ulp that then propagates - the difference is of the order of
    # the mix's own error (7e-3 wide / 2.5e-3 deep against fp32), and must stay well inside the logit tolerance
    assert err < 0.5 * LOGIT_ATOL, f"GPU mixed path differs from its emulation by {err:.3e}"


def test_mixed_config1_selection_identical_to_reference(mixed, golden):
    import b200al
    from oracle.synth import make_active, synth_pool_u8
    from oracle.unet_ref import OracleUNet
    cfg = np.load(os.path.join(ROOT, "tests", "golden", "configs.npz"))
    images = synth_pool_u8(32 + 512)
    model = OracleUNet(mixed[2][1])
    sel = b200al.B200EntropySelector(batch_size=12, num_workers=0, pin_memory=True, chunk_samples=64,
                                     precision=mixed["precision"])
    got = sel.select_next_batch(make_active(images, 32), 32, model, torch.device(DEV))
    assert got == cfg["c1/selected/entropy"].tolist()


def test_mixed_mc_dropout_and_embeddings(mixed, golden):
    from oracle.synth import synth_pool_u8
    eng, _sd = mixed[3]
    g = golden["mc_K3"]
    T, p, seed, start = int(g["T"]), float(g["p"]), int(g["seed"]), int(g["start"])
    imgs = synth_pool_u8(8, start=start).to(DEV)
    got = eng.scores(imgs, "entropy", T=T, drop_p=p, seed=seed, base_pos=0).cpu().numpy()
    rel = (np.abs(got - g["scores/entropy"]) / np.abs(g["scores/entropy"])).max()
    assert rel < 0.1 * SCORE_RTOL, f"MC-dropout scores: {rel:.3e}"
    # embeddings (k-center / k-means++ inputs) stay on the bf16x3 layer list: bit-identical to a bf16x3 engine
    import b200al
    ref = b200al.Engine(DEV, 3, max_samples=32)
    ref.load_model(_sd)
    assert torch.equal(ref.embed(imgs), eng.embed(imgs))
    assert torch.equal(ref.embed(imgs), eng.embed(imgs))   # ... also right after a mixed scoring pass
    ref.close()
