"""Wall-clock of the drop-in selectors on the other BASELINE.json configurations (1 GPU, or one process per
GPU under torchrun): config 1 (entropy, BUSI shape, 512-image pool, budget 32), config 3 (coreset-l2 k-center,
100k pool, 500 labeled, budget 500), config 4 (hybrid entropy -> k-center, 100k pool, budget 500).
Pools are synthetic (256 seeded base images tiled with circular shifts), resident in HBM as uint8.

Config "p": the input pipeline alone and through a selector - a pool of ragged FUGC-/BUSI-sized raw uint8 images
resident in HBM (device pool cache), resize + z-score statistics on the GPU per 128-image chunk.

    python tools/config_bench.py [--pool 100000] [--configs 1,3,4,p,k,b]
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "medical-image-analysis_b200")]

import numpy as np
import torch

import b200al
from oracle.prep_ref import synth_raw_u8
from oracle.synth import MirrorActive, MirrorExtendable, SynthBase, SynthRawBase, case_name, synth_pool_u8
from oracle.unet_ref import OracleUNet, golden_state_dict


def pool_u8(n, dev):
    base = synth_pool_u8(min(n, 256)).to(dev)
    reps = (n + base.shape[0] - 1) // base.shape[0]
    return torch.cat([torch.roll(base, shifts=(7 * r, 13 * r), dims=(1, 2)) for r in range(reps)])[:n].contiguous()


def active(images, n_labeled):
    names = [case_name(i) for i in range(images.shape[0])]
    base = SynthBase(images, names)
    return MirrorActive(MirrorExtendable(base, names[:n_labeled]), MirrorExtendable(base, names[n_labeled:]))


def timed(fn, reps=2):
    fn()  # warm-up (engine creation, weight packing)
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(reps):
        t0 = time.perf_counter()
        out = fn()
        torch.cuda.synchronize()
        best = min(best, time.perf_counter() - t0)
    return best, out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--pool", type=int, default=100_000)
    ap.add_argument("--configs", default="1,3,4")
    args = ap.parse_args()
    if "RANK" in os.environ:
        torch.distributed.init_process_group("nccl")
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    dev = torch.device("cuda", torch.cuda.current_device())
    g = np.load(os.path.join(ROOT, "tests", "golden", "model.npz"))
    kw = dict(batch_size=128, num_workers=0, pin_memory=False, chunk_samples=128)
    out = []
    for c in args.configs.split(","):
        if c == "1":
            model = OracleUNet(golden_state_dict(2, g))
            ad = active(pool_u8(512 + 32, dev), 32)
            sel = b200al.B200EntropySelector(**kw)
            dt, names = timed(lambda: sel.select_next_batch(ad, 32, model, dev))
            out.append({"config": 1, "what": "entropy, K=2, 512-image pool, budget 32", "seconds": dt,
                        "images_per_s": 512 / dt, "selected": len(names)})
        elif c in ("3", "3r"):
            # "3r": embeddings gathered, k-center replicated on every rank (the round-1 design) instead of the sharded
            # persistent kernel - same picks, for the stage comparison
            model = OracleUNet(golden_state_dict(3, g))
            ad = active(pool_u8(args.pool + 500, dev), 500)
            sel = b200al.B200CoresetSelector(metric="l2", shard_kcenter=(c == "3"), **kw)
            dt, names = timed(lambda: sel.select_next_batch(ad, 500, model, dev), reps=2)
            prof = b200al.B200CoresetSelector(metric="l2", shard_kcenter=(c == "3"), profile_stages=True, **kw)
            prof.select_next_batch(ad, 500, model, dev)
            import hashlib
            out.append({"config": c, "what": f"coreset l2 k-center, K=3, {args.pool}-image pool, 500 labeled, budget 500"
                                             + (" (k-center replicated)" if c == "3r" else ""),
                        "seconds": dt, "images_per_s": args.pool / dt, "selected": len(set(names)),
                        "picked_sha16": hashlib.sha256("\n".join(sorted(names)).encode()).hexdigest()[:16],
                        "stage_seconds_rank0": {k: round(v, 4) for k, v in prof.last_timing["stages"].items()}})
        elif c == "4":
            model = OracleUNet(golden_state_dict(3, g))
            ad = active(pool_u8(args.pool + 500, dev), 500)
            sel = b200al.B200HybridSelector(prefilter_factor=10, **kw)
            dt, names = timed(lambda: sel.select_next_batch(ad, 500, model, dev), reps=1)
            prof = b200al.B200HybridSelector(prefilter_factor=10, profile_stages=True, **kw)
            prof.select_next_batch(ad, 500, model, dev)
            out.append({"config": 4, "what": f"hybrid entropy top-5000 -> k-center, K=3, {args.pool}-image pool, budget 500",
                        "seconds": dt, "images_per_s": args.pool / dt, "selected": len(set(names)),
                        "stage_seconds_rank0": {k: round(v, 4) for k, v in prof.last_timing["stages"].items()}})
        elif c == "k":
            model = OracleUNet(golden_state_dict(3, g))
            ad = active(pool_u8(args.pool + 500, dev), 500)
            sel = b200al.B200KMeanSelector(metric="cosine", **kw)
            np.random.seed(0)
            dt, names = timed(lambda: sel.select_next_batch(ad, 500, model, dev), reps=1)
            out.append({"config": "k", "what": f"kmean-cosine (k-means++ seeding), K=3, {args.pool}-image pool, 500 labeled, budget 500",
                        "seconds": dt, "images_per_s": args.pool / dt, "selected": len(set(names))})
        elif c == "b":
            model = OracleUNet(golden_state_dict(3, g))
            n = min(args.pool, 10_000)
            ad = active(pool_u8(n + 32, dev), 32)
            sel = b200al.B200BADGESelector(**kw)
            np.random.seed(0)
            dt, names = timed(lambda: sel.select_next_batch(ad, 100, model, dev), reps=1)
            out.append({"config": "b", "what": f"badge (gradient embedding + k-means++), K=3, {n}-image pool, budget 100",
                        "seconds": dt, "images_per_s": n / dt, "selected": len(set(names))})
        elif c == "p":
            model = OracleUNet(golden_state_dict(3, g))
            n = min(args.pool, 10_000)
            distinct = [synth_raw_u8(i) for i in range(64)]
            raws = [distinct[i % 64] for i in range(n + 32)]
            base = SynthRawBase(raws, 256, normalize=True)
            ad = MirrorActive(MirrorExtendable(base, base.names[:32]), MirrorExtendable(base, base.names[32:]))
            sel = b200al.B200EntropySelector(**kw)
            dt, names = timed(lambda: sel.select_next_batch(ad, 32, model, dev))
            cache = base._b200_pool_cache
            eng = b200al.selectors.get_engine(dev, 3, 256, 256, "mixed", 128)
            rows = torch.arange(32, 32 + 1024, device=dev)
            for _ in range(2):
                cache.gather(eng, rows)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(5):
                cache.gather(eng, rows)
            e1.record()
            torch.cuda.synchronize()
            us = e0.elapsed_time(e1) * 1e3 / (5 * 1024)
            raw_b = sum(distinct[i % 64].size for i in range(1024)) / 1024
            out.append({"config": "p", "what": f"entropy with --do-normalize from the ragged raw pool cache, K=3, {n}-image pool",
                        "seconds": dt, "images_per_s": n / dt, "selected": len(names),
                        "prep_us_per_image": us, "prep_raw_bytes_per_image": raw_b,
                        "prep_GBps": (raw_b + 256 * 256 * 4) / us / 1e3, "cache_MB": cache.nbytes / 1e6})
    if rank == 0:
        for o in out:
            o["n_gpus"] = world
            print(json.dumps(o))


if __name__ == "__main__":
    main()
