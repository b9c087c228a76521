"""GPU bring-up diagnostics (not a test): per-layer tcgen05 vs SIMT errors, then
end-to-end numbers vs the golden vectors.  Usage: python tools/gpu_diag.py [layers|e2e|perf]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "medical-image-analysis_b200")]

import numpy as np
import torch

import b200al
from oracle.synth import synth_pool_u8
from oracle.unet_ref import golden_state_dict

G = {n: np.load(os.path.join(ROOT, "tests", "golden", f"{n}.npz")) for n in ("model", "pool_K2", "pool_K3", "mc_K3")}


def layers(precision):
    eng = b200al.Engine("cuda:0", 2, max_samples=2, precision=precision)
    eng.load_model(golden_state_dict(2, G["model"]))
    torch.cuda.synchronize()
    print(f"== layers precision={precision}", flush=True)
    for i in range(21):
        info = eng.debug_layer_info(i)
        g = torch.Generator(device="cuda").manual_seed(i)
        C = info["in_pitch"] // 2
        x = torch.randn((2, info["inH"], info["inW"], C), device="cuda", generator=g)
        hi = x.to(torch.bfloat16)
        lo = (x - hi.float()).to(torch.bfloat16)
        xin = torch.cat([hi, lo], dim=-1).contiguous()
        mask = (torch.rand((2, 2944), device="cuda", generator=g) > 0.3).float() / 0.7
        vals = []
        for impl in (1, 0):
            out = torch.zeros((2, info["outH"], info["outW"], info["out_pitch"]), device="cuda", dtype=torch.bfloat16)
            t0 = time.time()
            eng.debug_layer(i, impl, xin, out, mask)
            torch.cuda.synchronize()
            dt = time.time() - t0
            Ct = info["out_pitch"] // 2
            o = out.float()
            vals.append(((o[..., :Ct] + o[..., Ct:]) if precision == "bf16x3" else o[..., :Ct], dt))
        (ref, t_ref), (tc, t_tc) = vals
        err = (tc - ref).abs().max().item()
        scale = ref.abs().max().item()
        bad = (tc - ref).abs() > 1e-3 * scale
        print(f"layer {i:2d} kind={info['kind']} cin={info['cin']:3d} cout={info['cout']:3d} s={info['stride']} "
              f"out={info['outH']}x{info['outW']} err={err:.3e} scale={scale:.3e} rel={err / scale:.2e} "
              f"bad={int(bad.sum())}/{bad.numel()} t_tc={t_tc * 1e3:.2f}ms t_simt={t_ref * 1e3:.1f}ms", flush=True)
        if bad.any() and i < 21:
            idx = bad.nonzero()[:4].tolist()
            print("   first bad idx (s,y,x,c):", idx, "tc", [tc[tuple(j)].item() for j in idx],
                  "ref", [ref[tuple(j)].item() for j in idx], flush=True)
    eng.close()


def e2e(precision):
    print(f"== e2e precision={precision}", flush=True)
    for K, name in ((2, "pool_K2"), (3, "pool_K3")):
        g = G[name]
        L = int(g["n_labeled"])
        eng = b200al.Engine("cuda:0", K, max_samples=16, precision=precision)
        eng.load_model(golden_state_dict(K, G["model"]))
        imgs = synth_pool_u8(L + 16).to("cuda:0")
        for mode in ("entropy", "margin", "confidence"):
            s, lg = eng.scores(imgs[L:], mode, return_logits=True)
            want = g[f"scores/{mode}"][:16]
            rel = np.abs(s.cpu().numpy() - want) / np.abs(want)
            print(f"K={K} {mode}: max rel score err {rel.max():.3e}", flush=True)
        lerr = np.abs(lg[:2].cpu().numpy() - g["probe_logits"]).max()
        print(f"K={K} logits max abs err {lerr:.3e} (std {g['probe_logits'].std():.2f})", flush=True)
        emb = eng.embed(imgs).cpu().numpy()
        want = g["embeddings"][:L + 16]
        print(f"K={K} embedding max err / absmax {np.abs(emb - want).max() / np.abs(want).max():.3e}", flush=True)
        eng.close()


def perf(precision, chunk=64, n=256, T=1):
    print(f"== perf precision={precision} chunk={chunk} n={n} T={T}", flush=True)
    eng = b200al.Engine("cuda:0", 3, max_samples=chunk, precision=precision)
    eng.load_model(golden_state_dict(3, G["model"]))
    base = synth_pool_u8(32)
    imgs = base.repeat((n + 31) // 32, 1, 1)[:n].to("cuda:0")
    p = 0.1 if T > 1 else 0.0
    for _ in range(2):
        eng.scores(imgs, "entropy", T=T, drop_p=p, seed=1)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    eng.scores(imgs, "entropy", T=T, drop_p=p, seed=1)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    fl = b200al.conv_flops_per_sample() * n * T
    print(f"scores: {ms:.2f} ms for {n} images x T={T} -> {n / ms * 1e3:.0f} img/s, "
          f"{fl / ms / 1e9:.1f} TFLOP/s (algorithmic, whole pass)", flush=True)
    eng.set_timing(True)
    eng.scores(imgs, "entropy", T=T, drop_p=p, seed=1)
    cms, cn, cfl = eng.conv_time_ms()
    lt = eng.layer_times_ms()
    eng.set_timing(False)
    for i, ms_l in enumerate(lt):
        info = eng.debug_layer_info(i)
        px = info["outH"] * info["outW"] if info["kind"] == 0 else info["inH"] * info["inW"] * 4
        fl_l = 2.0 * px * info["cout"] * info["cin"] * (9 if info["kind"] == 0 else 1) * n * T
        print(f"  layer {i:2d} kind={info['kind']} {info['cin']:3d}->{info['cout']:3d} s{info['stride']} out {info['outH']:3d}: "
              f"{ms_l * 1e3 / (n * T):7.2f} us/sample {fl_l / ms_l / 1e9:6.1f} TF", flush=True)
    print(f"conv kernels: {cms:.2f} ms over {cn} launches, {cfl / cms / 1e9:.1f} TFLOP/s algorithmic", flush=True)
    e0.record()
    eng.embed(imgs)
    e1.record()
    torch.cuda.synchronize()
    print(f"embed: {e0.elapsed_time(e1):.2f} ms -> {n / e0.elapsed_time(e1) * 1e3:.0f} img/s", flush=True)
    eng.close()


if __name__ == "__main__":
    what = sys.argv[1] if len(sys.argv) > 1 else "layers"
    prec = sys.argv[2] if len(sys.argv) > 2 else "bf16x3"
    print(torch.cuda.get_device_name(0), flush=True)
    if what == "layers":
        layers(prec)
    elif what == "e2e":
        e2e(prec)
    elif what == "perf":
        perf(prec, *(int(a) for a in sys.argv[3:]))
