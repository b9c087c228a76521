"""Per-precision error table of the UNet forward on the golden probe images (VERDICT r01 item 3 / weak point 4).

  python tools/precision_table.py [--images 8] [--K 3] [--out profiles/r02_precision.json]

Emulates on the CPU, layer by layer, what each tensor-core operand format would do to the logits and to the entropy
score, against the fp32 oracle (oracle/unet_ref.py == the reference UNet, bit for bit):

  bf16     one bf16 operand each, fp32 accumulate                         1 MMA  per product
  fp16     one fp16 operand each                                          1 MMA
  tf32     10-bit-mantissa operands (kind::tf32 runs at half the bf16 rate: counted as 2)
  fp16x2   activations split hi+lo in fp16, weights single fp16           2 MMAs
  bf16x2   activations split hi+lo in bf16, weights single bf16           2 MMAs
  bf16x3   both operands split hi+lo in bf16, lo*lo dropped (the product path)   3 MMAs

"Whole net" rows run every conv / transposed conv in that format, with activations stored between layers the way the
format would store them (one value, or a hi/lo pair).  "Single layer" rows run ONE layer in the cheaper format and
all others in bf16x3 - the contribution a per-layer precision mix would have to budget for.  Tolerances of the north
star: logits 2e-2 absolute, scores 1e-3 relative.  This is an analysis tool (it uses the oracle as the fp32
reference); nothing on the product path imports it.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import numpy as np
import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]

from oracle import select_ref as S  # noqa: E402
from oracle.synth import synth_pool_u8  # noqa: E402
from oracle.unet_ref import BN_EPS, LEAKY_SLOPE, golden_state_dict, plain_blocks, unet_forward  # noqa: E402

MMAS = {"fp32": None, "bf16": 1, "fp16": 1, "tf32": 2, "fp16x2": 2, "bf16x2": 2, "bf16x3": 3}


def rnd(x: torch.Tensor, fmt: str) -> torch.Tensor:
    """x rounded to the operand format, held in fp32."""
    if fmt == "bf16":
        return x.to(torch.bfloat16).to(torch.float32)
    if fmt == "fp16":
        return x.to(torch.float16).to(torch.float32)
    if fmt == "tf32":  # round to nearest even at 10 mantissa bits
        i = x.contiguous().view(torch.int32)
        i = (i + 0x0FFF + ((i >> 13) & 1)) & ~0x1FFF
        return i.view(torch.float32)
    raise ValueError(fmt)


def split(x: torch.Tensor, fmt: str):
    hi = rnd(x, fmt)
    return hi, rnd(x - hi, fmt)


def contract(op, a: torch.Tensor, w: torch.Tensor, fmt: str) -> torch.Tensor:
    """op(a, w) (a bias-free conv / transposed conv) with operands in `fmt`, fp32 accumulation."""
    if fmt == "fp32":
        return op(a, w)
    if fmt in ("bf16", "fp16", "tf32"):
        return op(rnd(a, fmt), rnd(w, fmt))
    base = fmt[:-2]
    a_hi, a_lo = split(a, base)
    w_hi, w_lo = split(w, base)
    if fmt.endswith("x2"):
        return op(a_hi, w_hi) + op(a_lo, w_hi)
    return op(a_hi, w_hi) + op(a_hi, w_lo) + op(a_lo, w_hi)   # x3: lo*lo dropped


def store(x: torch.Tensor, fmt: str) -> torch.Tensor:
    """Activation as the next layer will read it back from HBM."""
    if fmt == "fp32":
        return x
    if fmt == "tf32":   # kind::tf32 reads fp32 words from shared memory: activations would stay fp32 in HBM
        return x
    if fmt in ("bf16", "fp16"):
        return rnd(x, fmt)
    hi, lo = split(x, fmt[:-2])
    return hi + lo


def forward(sd, x, fmts, storage):
    """UNet.forward with layer i (0..21 in launch order: 10 encoder convs, then per decoder level convT, conv, conv)
    contracted in fmts[i]; activations stored in `storage` format.  Layer 0 (Cin = 1, conv_first_kernel) and the 1x1
    head (fused epilogue) run in fp32 in the product path and stay fp32 here."""
    blocks = plain_blocks(1)
    li = 0
    if isinstance(storage, str):
        storage = [storage] * 22

    def block(i, z):
        nonlocal li
        prefix, _cin, _cout, stride = blocks[i]
        y = contract(lambda a, w: F.conv2d(a, w, None, stride=stride, padding=1), z, sd[f"{prefix}.0.weight"], fmts[li])
        li += 1
        y = y + sd[f"{prefix}.0.bias"][None, :, None, None]
        y = F.batch_norm(y, sd[f"{prefix}.2.running_mean"], sd[f"{prefix}.2.running_var"], sd[f"{prefix}.2.weight"],
                         sd[f"{prefix}.2.bias"], training=False, eps=BN_EPS)
        return store(F.leaky_relu(y, LEAKY_SLOPE), storage[li - 1])

    skips = []
    z = x
    for i in range(10):
        z = block(i, z)
        if i % 2 == 1:
            skips.append(z)
    y = skips[-1]
    for l in range(4):
        y = contract(lambda a, w: F.conv_transpose2d(a, w, None, stride=2), y, sd[f"decoder.upsamples.{l}.weight"], fmts[li])
        li += 1
        y = store(y + sd[f"decoder.upsamples.{l}.bias"][None, :, None, None], storage[li - 1])
        y = torch.cat([skips[3 - l], y], dim=1)
        y = block(10 + 2 * l, y)
        y = block(11 + 2 * l, y)
    assert li == 22
    return F.conv2d(y, sd["decoder.seg_output.weight"], sd["decoder.seg_output.bias"])


LAYER_NAMES = [f"encoder.levels.{l}.{j}" for l in range(5) for j in range(2)] + \
    [n for l in range(4) for n in (f"decoder.upsamples.{l}", f"decoder.levels.{l}.0", f"decoder.levels.{l}.1")]


# Per-layer precision mixes of the product path (layer indices as in LAYER_NAMES): the layers of `group` contract
# single fp16 operands (1 MMA, one plane); `entry` (bf16x3 itself) already stores its output as single fp16 for them,
# the last layer of the group stores a bf16 hi/lo pair again for the bf16x3 layers that follow.
MIXES = {
    "deep": {"entry": 6, "group": list(range(7, 14))},    # enc3.1 .. up1    (entry: enc3.0)
    "wide": {"entry": 4, "group": list(range(5, 17))},    # enc2.1 .. up2    (entry: enc2.0)
}


def mix_formats(name):
    m = MIXES[name]
    fmts = ["fp32"] + ["bf16x3"] * 21
    sto = ["bf16x3"] * 22
    for i in m["group"]:
        fmts[i] = "fp16"
        sto[i] = "fp16"
    sto[m["entry"]] = "fp16"
    sto[m["group"][-1]] = "bf16x3"
    return fmts, sto


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--images", type=int, default=8)
    ap.add_argument("--K", type=int, default=3)
    ap.add_argument("--threads", type=int, default=max(1, (os.cpu_count() or 2) // 2))
    ap.add_argument("--out", default=os.path.join(ROOT, "profiles", "r02_precision.json"))
    ap.add_argument("--no-layers", action="store_true")
    args = ap.parse_args()
    torch.set_num_threads(args.threads)
    sd = golden_state_dict(args.K, np.load(os.path.join(ROOT, "tests", "golden", "model.npz")))
    L = 16 if args.K == 3 else 32   # probe images = the first pool images of tests/golden/pool_K{K}.npz
    x = synth_pool_u8(args.images, start=L).float().div(255).unsqueeze(1)
    t0 = time.time()
    with torch.no_grad():
        ref = unet_forward(sd, x)
        ref_scores = S.entropy_scores(ref)
        harness = float((forward(sd, x, ["fp32"] * 22, "fp32") - ref).abs().max())
        assert harness < 1e-4, f"emulation harness differs from the oracle in fp32 by {harness}"  # bias added separately

        def measure(fmts, storage):
            lg = forward(sd, x, fmts, storage)
            sc = S.entropy_scores(lg)
            return {"max_abs_logit_err": float((lg - ref).abs().max()),
                    "max_rel_score_err": float(((sc - ref_scores).abs() / ref_scores.abs()).max())}

        out = {"model": f"golden model K={args.K} (tests/golden/model.npz)", "probe_images": args.images,
               "logit_std": float(ref.std()), "tolerance": {"logit_abs": 2e-2, "score_rel": 1e-3},
               "mma_per_product": {k: v for k, v in MMAS.items() if v},
               "fp32_harness_max_abs_logit_err": harness, "whole_net": {}, "single_layer": {}}
        for fmt in ("bf16", "fp16", "tf32", "bf16x2", "fp16x2", "bf16x3"):
            r = measure(["fp32"] + [fmt] * 21, fmt)
            r["meets_tolerance"] = r["max_abs_logit_err"] < 2e-2 and r["max_rel_score_err"] < 1e-3
            out["whole_net"][fmt] = r
            print(f"whole net {fmt:7s} logit {r['max_abs_logit_err']:.3e} score {r['max_rel_score_err']:.3e} "
                  f"{'ok' if r['meets_tolerance'] else 'FAILS'}  ({time.time() - t0:.0f}s)", flush=True)
        out["mixes"] = {}
        for name in MIXES:
            fm, sto = mix_formats(name)
            r = measure(fm, sto)
            r["fp16_layers"] = [LAYER_NAMES[i] for i in MIXES[name]["group"]]
            r["entry_layer"] = LAYER_NAMES[MIXES[name]["entry"]]
            out["mixes"][name] = r
            print(f"mix {name:5s} logit {r['max_abs_logit_err']:.3e} score {r['max_rel_score_err']:.3e}", flush=True)
        if not args.no_layers:
            for fmt in ("fp16", "fp16x2", "bf16"):
                rows = {}
                for i, name in enumerate(LAYER_NAMES):
                    if i == 0:   # encoder.levels.0.0 (Cin = 1) runs in fp32 on CUDA cores in the product path
                        continue
                    fm = ["fp32"] + ["bf16x3"] * 21
                    fm[i] = fmt
                    rows[name] = measure(fm, "bf16x3")
                    print(f"layer {name:22s} in {fmt:7s} logit {rows[name]['max_abs_logit_err']:.3e} "
                          f"score {rows[name]['max_rel_score_err']:.3e}", flush=True)
                out["single_layer"][fmt] = rows
    out["seconds"] = round(time.time() - t0, 1)
    with open(args.out, "w") as f:
        json.dump(out, f, indent=1)
    print("wrote", args.out)


if __name__ == "__main__":
    main()
