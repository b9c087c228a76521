"""tcgen05 implicit-GEMM conv / transposed conv vs the SIMT cross-check kernel
on the same packed operands, every layer shape of the UNet (GPU only)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def engine(golden):
    import b200al
    from oracle.unet_ref import golden_state_dict
    eng = b200al.Engine("cuda:0", 2, max_samples=4)
    eng.load_model(golden_state_dict(2, golden["model"]))
    yield eng
    eng.close()


def _run_layer(eng, i, S=2, with_mask=False, seed=0):
    info = eng.debug_layer_info(i)
    g = torch.Generator(device="cuda").manual_seed(seed + i)
    C = info["in_pitch"] // 2
    x = torch.randn((S, info["inH"], info["inW"], C), device="cuda", generator=g)
    hi = x.to(torch.bfloat16)
    lo = (x - hi.float()).to(torch.bfloat16)
    xin = torch.cat([hi, lo], dim=-1).contiguous()
    mask = None
    if with_mask:
        mask = (torch.rand((S, 2944), device="cuda", generator=g) > 0.3).float() / 0.7
    outs = []
    for impl in (0, 1):
        out = torch.zeros((S, info["outH"], info["outW"], info["out_pitch"]), device="cuda", dtype=torch.bfloat16)
        eng.debug_layer(i, impl, xin, out, mask)
        torch.cuda.synchronize()
        Ct = info["out_pitch"] // 2
        o = out.float()
        val = o[..., :Ct] + o[..., Ct:]
        outs.append((val, out))
    return info, outs


@pytest.mark.parametrize("layer", list(range(21)))
def test_layer_matches_simt(engine, layer):
    info, ((tc, tc_raw), (ref, ref_raw)) = _run_layer(engine, layer, with_mask=(layer % 2 == 0))
    c0, c1 = info["out_c_off"], info["out_c_off"] + info["cout"]
    assert ref[..., c0:c1].abs().max() > 0
    err = (tc - ref).abs().max().item()
    scale = ref.abs().max().item()
    assert err <= 2e-4 * scale, f"layer {layer} {info}: max err {err} vs scale {scale}"
    # channels outside the layer's output window stay untouched
    Ct = info["out_pitch"] // 2
    untouched = torch.ones(Ct, dtype=torch.bool)
    untouched[c0:c1] = False
    assert tc_raw[..., :Ct][..., untouched].abs().max().item() == 0 if untouched.any() else True
