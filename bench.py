"""Acquisition-pass benchmark: pool images/s scored + selected.

  python bench.py --gpus N --steps K --warmup W            (our arm)
  python bench.py --impl reference --gpus N --steps K ...  (CPU reference arm)

Workload at every N = BASELINE.json configs[1]: UNet MC-dropout (T=8) entropy acquisition on ONE FUGC-shape
(1x256x256, K=3) synthetic pool of 10 000 images, budget 100, through the public drop-in call
`B200MCDropoutEntropySelector.select_next_batch`.  For N > 1 the ranks run the same call under the NCCL process
group: the pool is sharded into contiguous position ranges (STRONG scaling - the total work is fixed), every rank
scores its shard, and only the per-rank top-k candidates are all-gathered and merged.  One step = one full
`select_next_batch`: weight broadcast + fold/pack, MC masks, forward, uncertainty reduction, rank & cut, merge,
position -> name mapping.

value  : images/s with the uint8 pool already resident in HBM (the dataset's `images_u8` is a CUDA tensor).
e2e    : the same call with the pool in pinned HOST memory: per-chunk H2D copies of the rank's shard and the D2H
         read of the selected positions are inside the timed region.
roofline: the tcgen05 conv kernels (all 21 layers), CUDA-event timed per launch inside the library during the
         timed steps of the `value` arm.
config4: secondary record - BASELINE.json configs[3], hybrid entropy -> k-center over a 100k-image pool, budget 500,
         through `B200HybridSelector.select_next_batch` (seconds, images/s, stage breakdown, hash of the picks).
cpu_baseline: the oracle port (torch fp32 restatement of the reference path) on the host cores, bounded sample.
"""
from __future__ import annotations

import argparse
import hashlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path[:0] = [ROOT, os.path.join(ROOT, "medical-image-analysis_b200")]

POOL = 10_000
T_REPEATS = 8
DROP_P = 0.1
BUDGET = 100
K_CLASSES = 3
MC_SEED = 2025
METRIC = "pool images/sec scored+selected"
C4_POOL, C4_LABELED, C4_BUDGET = 100_000, 500, 500


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--pool", type=int, default=POOL, help="images in the (one, sharded) pool")
    ap.add_argument("--chunk", type=int, default=int(os.environ.get("B200AL_CHUNK", 256)),
                    help="samples per forward chunk (256 = 20 GB of activations)")
    ap.add_argument("--precision", default="mixed", choices=["mixed", "mixed-deep", "bf16x3", "bf16"])
    ap.add_argument("--cpu-sample", type=int, default=12, help="images in the CPU-baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-config4", action="store_true", help="skip the secondary hybrid 100k record")
    ap.add_argument("--c4-pool", type=int, default=C4_POOL)
    return ap.parse_args()


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return p.get("bf16_tflops_sustained", 1400.0), p.get("hbm_gbs", 6650.0), "measured (MEASURED_PEAKS.json, sustained)"
    return 1400.0, 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""

    def __init__(self, index: int):
        self.rows, self.proc = [], None
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={index}", f"--query-gpu={q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i].lower() == "active" for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


DTYPES = {
    "bf16x3": "bf16x3 (bf16 hi/lo split operands, 3 MMAs per product, fp32 accumulate)",
    "bf16": "bf16",
    "mixed": "per-layer mix, fp32 accumulate: fp16 single operands (1 MMA) in encoder.levels.2.1 ... decoder.upsamples.2 "
             "(12 layers, <= 64x64), bf16x3 in the other 9 tensor-core layers, fp32 first conv / head",
    "mixed-deep": "per-layer mix, fp32 accumulate: fp16 single operands (1 MMA) in encoder.levels.3.1 ... "
                  "decoder.upsamples.1 (7 layers, <= 32x32), bf16x3 in the other 14 tensor-core layers",
}


def issue_factor(precision: str) -> float:
    """MMAs issued per algorithmic MAC, FLOP-weighted over the 21 tensor-core layers (SURVEY 8a' per-layer GFLOP)."""
    if precision == "bf16":
        return 1.0
    gf = [1.208, 0.604, 1.208, 0.604, 1.208, 0.604, 1.208, 0.604, 1.208, 0.268, 2.416, 1.208, 0.268, 2.416, 1.208,
          0.268, 2.416, 1.208, 0.268, 2.416, 1.208]   # library layer order (first conv excluded)
    group = {"bf16x3": range(0), "mixed": range(4, 16), "mixed-deep": range(6, 13)}[precision]
    return sum(g * (1 if i in group else 3) for i, g in enumerate(gf)) / sum(gf)


def workload_name(args):
    return (f"UNet MC-dropout (T={T_REPEATS}) entropy acquisition, FUGC-shape 1x256x256 K={K_CLASSES}, "
            f"{args.pool}-image synthetic pool of distinct images (one pool, sharded across the GPUs), budget {BUDGET}")


def names_hash(names) -> str:
    return hashlib.sha256("\n".join(map(str, names)).encode()).hexdigest()[:16]


# ---------------------------------------------------------------------------------------------------------------
# CPU reference arm / cpu_baseline: the oracle port of the reference path (the only place bench.py executes oracle/)
# ---------------------------------------------------------------------------------------------------------------
def cpu_pass(sd, images_u8, positions):
    """One bounded sample of the same workload on the host: composed MC-dropout oracle (reference UNet arithmetic,
    explicit masks, reference entropy formula) + stable rank."""
    from oracle import select_ref as S
    scores = S.pool_scores_mc(sd, images_u8, positions, T_REPEATS, DROP_P, MC_SEED, "entropy", batch=4)
    return S.rank_desc_stable(scores, min(BUDGET, len(scores)))


def cpu_baseline(sample: int):
    import numpy as np
    import torch
    from bench_fixtures import seeded_state_dict, synthetic_pool_u8
    torch.set_num_threads(os.cpu_count())
    sd = seeded_state_dict(K_CLASSES)
    imgs = synthetic_pool_u8(sample)
    cpu_pass(sd, imgs[:2], np.arange(2))  # warm up thread pools / allocator
    t0 = time.perf_counter()
    cpu_pass(sd, imgs, np.arange(sample))
    dt = time.perf_counter() - t0
    return {"value": sample / dt, "unit": "images/s", "cores": os.cpu_count(), "kind": "port",
            "sample": f"{sample} images x T={T_REPEATS} MC-dropout forward + entropy + rank (oracle/select_ref.pool_scores_mc), "
                      f"torch {torch.__version__} fp32, {torch.get_num_threads()} threads"}


def run_reference(args):
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    import numpy as np
    import torch
    from bench_fixtures import seeded_state_dict, synthetic_pool_u8
    torch.set_num_threads(os.cpu_count())
    sd = seeded_state_dict(K_CLASSES)
    per_step = max(2, args.cpu_sample // 3)
    imgs = synthetic_pool_u8(per_step)
    for _ in range(min(args.warmup, 1)):
        cpu_pass(sd, imgs[:2], np.arange(2))
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_pass(sd, imgs, np.arange(per_step))
    dt = time.perf_counter() - t0
    v = per_step * args.steps / dt
    cb = {"value": v, "unit": "images/s", "cores": os.cpu_count(), "kind": "port",
          "sample": f"{per_step} images x T={T_REPEATS} per step (oracle port of the reference path, torch fp32 CPU; "
                    "MC-dropout does not exist in the reference, so the port composes its UNet + Dropout2d + entropy)"}
    emit({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "images/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(args), "note": "CPU host cores; each step is a bounded sample"},
        "cpu_baseline": cb,
        "e2e": {"value": v, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    })


def ncu_traffic():
    """DRAM bytes per launch (read + write) from the COMMITTED `ncu --set full` summaries of the same kernels
    (static: not measured by this run).  conv: mean over the conv launches of one 128-sample chunk; uncertainty: per
    image.  None when the files are absent."""
    import csv

    def rows(name):
        path = os.path.join(ROOT, "profiles", name)
        if not os.path.exists(path):
            return []
        with open(path) as f:
            return list(csv.DictReader(f))

    def total(r):
        out = 0.0
        for k, v in r.items():
            if k.startswith("dram__bytes_read.sum") or k.startswith("dram__bytes_write.sum"):
                unit = k[k.index("[") + 1:k.index("]")]
                out += float(v) * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}[unit]
        return out

    conv_file = next((f for f in ("r02_ncu_full_conv.csv", "r01_ncu_full_conv_v11.csv") if rows(f)), None)
    unc_file, unc_imgs = next(((f, n) for f, n in (("r02_ncu_full_uncertainty_128.csv", 128),
                                                   ("r01_ncu_full_uncertainty_v9.csv", 16)) if rows(f)), (None, 1))
    conv = [total(r) for r in rows(conv_file)] if conv_file else []
    unc = [total(r) for r in rows(unc_file)] if unc_file else []
    return ((sum(conv) / len(conv) if conv else None), conv_file,
            (sum(unc) / len(unc) / unc_imgs if unc else None), unc_file)


# ---------------------------------------------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------------------------------------------
def run_b200(args):
    import torch
    import torch.distributed as dist

    import b200al
    from b200al import selectors as SM
    from bench_fixtures import StateDictModel, make_active, seeded_state_dict, synthetic_pool_u8

    world = int(os.environ.get("WORLD_SIZE", 1))
    rank = int(os.environ.get("RANK", 0))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    n = args.pool
    model = StateDictModel(seeded_state_dict(K_CLASSES), dropout_p=DROP_P)

    # ONE pool (every rank builds the same one): image 0 is the labeled set, images 1..n the pool.
    # The selector shards pool positions by contiguous ranges (b200al.dist.shard_range) and reads only its shard.
    # every pool image is a DISTINCT seeded image (tiled near-duplicates would put near-ties at the cut by construction)
    host = synthetic_pool_u8(n + 1, distinct=n + 1).pin_memory()
    ad_resident = make_active(host.to(dev), 1)
    ad_host = make_active(host, 1)
    lo, hi = b200al.dist.shard_range(n, rank, world)
    sel = b200al.B200MCDropoutEntropySelector(batch_size=args.chunk, num_workers=0, pin_memory=True,
                                              mc_repeats=T_REPEATS, dropout_prob=DROP_P, mc_seed=MC_SEED,
                                              precision=args.precision, chunk_samples=args.chunk)
    eng = SM.get_engine(dev, K_CLASSES, 256, 256, args.precision, args.chunk)  # the engine the selectors will use

    def step_resident():
        return sel.select_next_batch(ad_resident, BUDGET, model, dev)

    def step_e2e():
        return sel.select_next_batch(ad_host, BUDGET, model, dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            out = fn()
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, out

    for _ in range(args.warmup):
        picked = step_resident()
    barrier()
    launches0 = eng.kernel_launches()
    sampler = ClockSampler(local) if rank == 0 else None
    eng.set_timing(True)
    ms, picked = timed(step_resident, args.steps)
    conv_ms, conv_n, conv_fl = eng.conv_time_ms()
    eng.set_timing(False)
    clocks = sampler.stop() if sampler else None
    launches = eng.kernel_launches() - launches0
    if world > 1:
        t = torch.tensor([float(launches)], device=dev, dtype=torch.float64)
        dist.all_reduce(t)
        launches = int(t.item())

    e2e_steps = 5 if args.steps >= 5 else max(1, args.steps)
    step_e2e()  # warm the pinned staging path
    ms_e2e, picked_e2e = timed(step_e2e, e2e_steps)

    # secondary record: BASELINE config 4 (hybrid entropy -> k-center, 100k pool, budget 500) through the public API
    c4 = None
    if not args.no_config4:
        c4 = run_config4(args, dev, rank, world, timed)

    # kernel 3 of the north star on its own: softmax + entropy + mean over fp32 logits resident in HBM
    # (805 MB per launch, larger than L2); algorithmic bytes = T*K*H*W*4 + 4 per image
    unc = None
    if rank == 0:
        n_u = 128
        lg = torch.randn((n_u, T_REPEATS, K_CLASSES, 256, 256), device=dev)
        for _ in range(3):
            eng.uncertainty_from_logits(lg, "entropy")
        torch.cuda.synchronize()
        u0, u1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        u0.record()
        for _ in range(5):
            eng.uncertainty_from_logits(lg, "entropy")
        u1.record()
        torch.cuda.synchronize()
        unc_ms = u0.elapsed_time(u1) / 5
        unc_bytes = n_u * (T_REPEATS * K_CLASSES * 256 * 256 * 4 + 4)
        unc = (unc_bytes / (unc_ms / 1e3) / 1e9, unc_ms, unc_bytes)
        del lg

    if rank == 0:
        tf_peak, hbm_peak, peak_src = peaks()
        value = n * args.steps / (ms / 1e3)
        achieved = conv_fl / (conv_ms / 1e3) / 1e12 if conv_ms > 0 else 0.0
        conv_traffic, conv_file, unc_traffic_per_image, unc_file = ncu_traffic()
        line = {
            "metric": METRIC, "value": value, "unit": "images/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": DTYPES[args.precision], "data": "synthetic",
            "config": {"workload": workload_name(args), "chunk_samples": args.chunk,
                       "api": "B200MCDropoutEntropySelector.select_next_batch (value: pool resident in HBM; e2e: pool in "
                              "pinned host memory)",
                       "sharding": f"contiguous position ranges, {hi - lo} images on rank 0",
                       "l2": "inputs larger than L2 (655 MB pool, >5 GB activations per chunk pass)",
                       "parity": "MC-dropout is a composed oracle: parity unpinned by the reference (SURVEY F6); "
                                 "checked against it at T=8 on the 1000-image prefix (tests/test_gpu_configs.py)",
                       "selection_consistent_resident_vs_e2e": list(picked) == list(picked_e2e),
                       "picked_set_sha16": names_hash(sorted(picked))},
            "e2e": {"value": n * e2e_steps / (ms_e2e / 1e3), "unit": "images/s", "steps": e2e_steps,
                    "h2d_bytes_per_step": n * 256 * 256, "d2h_bytes_per_step": BUDGET * 12 * world},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {"bound": "tensor", "achieved": achieved, "peak": tf_peak, "unit": "TFLOP/s",
                         "frac": achieved / tf_peak,
                         "traffic": conv_traffic * args.chunk / 128 if conv_traffic else None,
                         "traffic_note": f"static, from profiles/{conv_file}: mean DRAM read+write bytes per conv launch of "
                                         "an `ncu --set full` capture of a 128-sample chunk, scaled to this run's "
                                         "chunk_samples (traffic is per sample); not measured by this run",
                         "kernel": "conv_fold_kernel + conv_tc_kernel (tcgen05 implicit GEMM, the 21 conv/convT layers), "
                                   "rank 0's launches",
                         "launches_timed": int(conv_n), "avg_launch_ms": conv_ms / max(conv_n, 1),
                         "flops_note": "algorithmic 2*MAC of the fp32 reference conv; bf16x3 issues 3 MMAs per product, "
                                       "so the tensor pipe runs at >= 3x this rate (frac_of_issue_ceiling = achieved / (peak/3))",
                         "frac_of_issue_ceiling": achieved * issue_factor(args.precision) / tf_peak,
                         "peak_source": peak_src},
        }
        if unc:
            line["roofline_hbm_kernel"] = {"bound": "hbm", "achieved": unc[0], "peak": hbm_peak, "unit": "GB/s",
                                           "frac": unc[0] / hbm_peak,
                                           "traffic": unc_traffic_per_image * 128 if unc_traffic_per_image else None,
                                           "traffic_note": f"static, from profiles/{unc_file} (ncu --set full DRAM read+write "
                                                           "bytes per image x 128 images); algorithmic = bytes_per_launch",
                                           "kernel": "uncertainty_logits_kernel (softmax + entropy + per-image mean, "
                                                     "128 images x T=8 x K=3 fp32 logits)",
                                           "avg_launch_ms": unc[1], "bytes_per_launch": unc[2],
                                           "peak_source": peak_src}
        if c4:
            line["config4"] = c4
        if not args.no_cpu_baseline and world == 1:
            line["cpu_baseline"] = cpu_baseline(args.cpu_sample)
        emit(line)
    if world > 1:
        dist.destroy_process_group()


def run_config4(args, dev, rank, world, timed):
    """BASELINE.json configs[3]: hybrid uncertainty + coreset over ONE 100k-image FUGC-shape pool (500 labeled, budget
    500, entropy pre-filter to 10 x budget), pool resident in HBM, sharded by the selector under the process group."""
    import torch

    import b200al
    from bench_fixtures import StateDictModel, make_active, seeded_state_dict, synthetic_pool_u8
    n = args.c4_pool
    model = StateDictModel(seeded_state_dict(K_CLASSES))
    images = synthetic_pool_u8(C4_LABELED + n)
    ad = make_active(images.to(dev), C4_LABELED)
    del images
    kw = dict(batch_size=args.chunk, num_workers=0, pin_memory=True, precision=args.precision, chunk_samples=args.chunk)
    sel = b200al.B200HybridSelector(**kw)
    picked = sel.select_next_batch(ad, C4_BUDGET, model, dev)   # warm
    ms, picked = timed(lambda: sel.select_next_batch(ad, C4_BUDGET, model, dev), 2)
    prof = b200al.B200HybridSelector(profile_stages=True, **kw)
    prof.select_next_batch(ad, C4_BUDGET, model, dev)
    stages = {k: round(v, 4) for k, v in prof.last_timing["stages"].items()}
    del ad
    torch.cuda.empty_cache()
    return {"workload": f"hybrid entropy -> k-center (l2), FUGC-shape, {n}-image pool (one pool, sharded), "
                        f"{C4_LABELED} labeled, budget {C4_BUDGET}, B200HybridSelector.select_next_batch",
            "seconds": ms / 2 / 1e3, "images_per_s": n / (ms / 2 / 1e3), "steps": 2,
            "picked_sha16": names_hash(sorted(picked)), "n_picked": len(picked),
            "stage_seconds_rank0": stages,
            "stage_note": "separate profiled call (device synchronised at stage boundaries)"}


def _claim_stdout():
    """stdout carries exactly ONE JSON line, but libraries write there too (NCCL prints its version banner to fd 1
    whatever NCCL_DEBUG_FILE says).  Keep the real stdout for the result and point fd 1 at stderr for everyone else."""
    global _RESULT_OUT
    sys.stdout.flush()
    _RESULT_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)


def emit(line: dict):
    _RESULT_OUT.write(json.dumps(line) + "\n")
    _RESULT_OUT.flush()


if __name__ == "__main__":
    a = parse()
    _claim_stdout()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
