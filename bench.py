"""Acquisition-pass benchmark: pool images/s scored + selected.

  python bench.py --gpus N --steps K --warmup W            (our arm)
  python bench.py --impl reference --gpus N --steps K ...  (CPU reference arm)

Workload at every N = BASELINE.json configs[1]: UNet MC-dropout (T=8) entropy
acquisition on a FUGC-shape (1x256x256, K=3) synthetic pool, 10 000 images per
GPU, budget 100.  One step = one full acquisition pass over the rank's pool
(score every image, rank, cut) plus, for N > 1, the all-gather + merge of the
per-rank top-k candidates.  Weak scaling: each rank owns its own 10k shard.

value  : images/s with the uint8 pool already resident in HBM.
e2e    : the same pass through the drop-in selector (`select_next_batch`) with
         the pool in pinned HOST memory: per-chunk H2D copies and the D2H read
         of the selected positions are inside the timed region.
roofline: the tcgen05 conv kernels (all layers), CUDA-event timed per launch
         inside the library during the timed steps.
cpu_baseline: the oracle port (torch fp32 restatement of the reference path) on
         the host cores, on a bounded sample.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path[:0] = [ROOT, os.path.join(ROOT, "medical-image-analysis_b200")]

POOL_PER_GPU = 10_000
T_REPEATS = 8
DROP_P = 0.1
BUDGET = 100
K_CLASSES = 3
MC_SEED = 2025
METRIC = "pool images/sec scored+selected"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--pool", type=int, default=POOL_PER_GPU, help="images per GPU")
    ap.add_argument("--chunk", type=int, default=int(os.environ.get("B200AL_CHUNK", 256)),
                    help="samples per forward chunk (256 = 20 GB of activations; 128 -> 256 is +1.8 %, 512 another +0.4 %)")
    ap.add_argument("--precision", default="bf16x3", choices=["bf16x3", "bf16"])
    ap.add_argument("--cpu-sample", type=int, default=12, help="images in the CPU-baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
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


def golden_model():
    import numpy as np
    from oracle.unet_ref import golden_state_dict
    g = np.load(os.path.join(ROOT, "tests", "golden", "model.npz"))
    return golden_state_dict(K_CLASSES, g)


def synthetic_pool_u8(n: int, rank: int):
    """[n,256,256] uint8 on the CPU: 256 seeded base images (oracle.synth),
    tiled with per-copy circular shifts so every pool image is distinct."""
    import torch
    from oracle.synth import synth_pool_u8
    base = synth_pool_u8(min(n, 256), start=rank * 1_000_000)
    reps = (n + base.shape[0] - 1) // base.shape[0]
    parts = [torch.roll(base, shifts=(7 * r, 13 * r), dims=(1, 2)) for r in range(reps)]
    return torch.cat(parts)[:n].contiguous()


# ---------------------------------------------------------------------------
# CPU reference arm / cpu_baseline
# ---------------------------------------------------------------------------
def cpu_pass(sd, images_u8, positions):
    """One bounded sample of the same workload on the host: composed MC-dropout
    oracle (reference UNet arithmetic, explicit masks, reference entropy
    formula) + stable rank."""
    from oracle import select_ref as S
    scores = S.pool_scores_mc(sd, images_u8, positions, T_REPEATS, DROP_P, MC_SEED, "entropy", batch=4)
    return S.rank_desc_stable(scores, min(BUDGET, len(scores)))


def cpu_baseline(sample: int):
    import numpy as np
    import torch
    torch.set_num_threads(os.cpu_count())
    sd = golden_model()
    imgs = synthetic_pool_u8(sample, 0)
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
    torch.set_num_threads(os.cpu_count())
    sd = golden_model()
    per_step = max(2, args.cpu_sample // 3)
    imgs = synthetic_pool_u8(per_step, 0)
    for _ in range(min(args.warmup, 1)):
        cpu_pass(sd, imgs[:2], np.arange(2))
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_pass(sd, imgs, np.arange(per_step))
    dt = time.perf_counter() - t0
    v = per_step * args.steps / dt
    cb = {"value": v, "unit": "images/s", "cores": os.cpu_count(), "kind": "port",
          "sample": f"{per_step} images x T={T_REPEATS} per step (oracle port of the reference path, torch fp32 CPU)"}
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "images/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(args), "note": "CPU host cores; each step is a bounded sample"},
        "cpu_baseline": cb,
        "e2e": {"value": v, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def ncu_traffic():
    """DRAM bytes per launch (read + write) from the committed `ncu --set full` summaries of the same kernels at the
    bench's chunk size (profiles/r01_ncu_full_conv_v11.csv: the conv launches of one 128-sample chunk;
    profiles/r01_ncu_full_uncertainty_v9.csv: 16 images x T=8 per launch).  None when the files are absent."""
    import csv
    here = os.path.dirname(os.path.abspath(__file__))

    def rows(name):
        path = os.path.join(here, "profiles", name)
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

    conv = [total(r) for r in rows("r01_ncu_full_conv_v11.csv")]
    unc = [total(r) for r in rows("r01_ncu_full_uncertainty_v9.csv")]
    return (sum(conv) / len(conv) if conv else None), (sum(unc) / len(unc) / 16 if unc else None)


def workload_name(args):
    return (f"UNet MC-dropout (T={T_REPEATS}) entropy acquisition, FUGC-shape 1x256x256 K={K_CLASSES}, "
            f"{args.pool}-image synthetic pool per GPU, budget {BUDGET}")


# ---------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------
def run_b200(args):
    import torch
    import torch.distributed as dist

    import b200al
    from b200al import dist as D
    from oracle.synth import MirrorActive, MirrorExtendable, SynthBase, case_name
    from oracle.unet_ref import OracleUNet

    world = int(os.environ.get("WORLD_SIZE", 1))
    rank = int(os.environ.get("RANK", 0))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    n = args.pool
    sd = golden_model()
    model = OracleUNet(sd)
    pool_cpu = synthetic_pool_u8(n, rank).pin_memory()
    pool_dev = pool_cpu.to(dev)

    eng = b200al.Engine(dev, K_CLASSES, max_samples=args.chunk, precision=args.precision)
    eng.load_model(model)
    lo = rank * n  # global positions of this rank's shard (weak scaling: shards are disjoint pools)

    def step_resident():
        # MC-dropout masks are keyed by the position inside the rank's own pool (as in the selector path below)
        scores = eng.scores(pool_dev, "entropy", T=T_REPEATS, drop_p=DROP_P, seed=MC_SEED, base_pos=0)
        vals, pos = eng.topk(scores, BUDGET, base_pos=lo)
        if world > 1:
            vals, pos = D.global_topk(vals, pos, BUDGET)
        return pos

    # e2e: the public selector API over a host-resident pool (this rank's shard is its whole pool)
    names = [case_name(lo + i) for i in range(n)] + ["labeled-0"]
    base = SynthBase(torch.cat([pool_cpu, pool_cpu[:1]]), names)
    base.images_u8 = base.images_u8.pin_memory()
    ad = MirrorActive(MirrorExtendable(base, names[n:]), MirrorExtendable(base, names[:n]))
    sel = b200al.B200MCDropoutEntropySelector(batch_size=args.chunk, num_workers=0, pin_memory=True,
                                              mc_repeats=T_REPEATS, dropout_prob=DROP_P, mc_seed=MC_SEED,
                                              precision=args.precision, chunk_samples=args.chunk)
    sel_engine = {}

    def step_e2e():
        # single-process semantics per rank (each rank's pool is its own); merge candidates afterwards
        e, scores, _lo, _hi = _scores_whole_pool(sel, ad, model, dev)
        sel_engine["e"] = e
        vals, pos = e.topk(scores, BUDGET, base_pos=lo)
        if world > 1:
            vals, pos = D.global_topk(vals, pos, BUDGET)
        return pos.cpu()  # D2H read of the result

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

    step_e2e()  # warm the selector's engine / pinned staging
    ms_e2e, picked_e2e = timed(step_e2e, max(1, min(args.steps, 2)))
    e2e_steps = max(1, min(args.steps, 2))

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

    same = sorted(picked.cpu().tolist()) == sorted(picked_e2e.tolist())
    if rank == 0:
        tf_peak, hbm_peak, peak_src = peaks()
        value = world * n * args.steps / (ms / 1e3)
        achieved = conv_fl / (conv_ms / 1e3) / 1e12 if conv_ms > 0 else 0.0
        conv_traffic, unc_traffic_per_image = ncu_traffic()
        line = {
            "metric": METRIC, "value": value, "unit": "images/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16x3 (bf16 hi/lo split operands, fp32 accumulate)"
            if args.precision == "bf16x3" else "bf16", "data": "synthetic",
            "config": {"workload": workload_name(args), "chunk_samples": args.chunk,
                       "l2": "inputs larger than L2 (655 MB pool, >5 GB activations per chunk pass)",
                       "selection_consistent_resident_vs_e2e": same},
            "e2e": {"value": world * n * e2e_steps / (ms_e2e / 1e3), "unit": "images/s",
                    "h2d_bytes_per_step": n * 256 * 256, "d2h_bytes_per_step": BUDGET * 8},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {"bound": "tensor", "achieved": achieved, "peak": tf_peak, "unit": "TFLOP/s",
                         "frac": achieved / tf_peak,
                         "traffic": conv_traffic * args.chunk / 128 if conv_traffic else None,
                         "traffic_note": "mean DRAM read+write bytes per conv launch, ncu --set full of a 128-sample chunk "
                                         "(profiles/r01_ncu_full_conv_v11.csv) scaled to this run's chunk_samples "
                                         "(traffic is per sample: activations in, activations out); tensor-bound "
                                         "kernels, shown for completeness",
                         "kernel": "conv_fold_kernel + conv_tc_kernel (tcgen05 implicit GEMM, the 21 conv/convT layers)",
                         "launches_timed": int(conv_n), "avg_launch_ms": conv_ms / max(conv_n, 1),
                         "flops_note": "algorithmic 2*MAC of the fp32 reference conv; bf16x3 issues 3 MMAs per product, "
                                       "so the tensor pipe runs at >= 3x this rate (frac_of_issue_ceiling = achieved / (peak/3))",
                         "frac_of_issue_ceiling": (achieved * (3 if args.precision == "bf16x3" else 1)) / tf_peak,
                         "peak_source": peak_src},
        }
        if unc:
            line["roofline_hbm_kernel"] = {"bound": "hbm", "achieved": unc[0], "peak": hbm_peak, "unit": "GB/s",
                                           "frac": unc[0] / hbm_peak,
                                           "traffic": unc_traffic_per_image * 128 if unc_traffic_per_image else None,
                                           "traffic_note": "ncu --set full DRAM read+write bytes per image x 128 images "
                                                           "(profiles/r01_ncu_full_uncertainty_v9.csv); algorithmic = "
                                                           "bytes_per_launch",
                                           "kernel": "uncertainty_logits_kernel (softmax + entropy + per-image mean, "
                                                     "128 images x T=8 x K=3 fp32 logits)",
                                           "avg_launch_ms": unc[1], "bytes_per_launch": unc[2],
                                           "peak_source": peak_src}
        if not args.no_cpu_baseline and world == 1:
            line["cpu_baseline"] = cpu_baseline(args.cpu_sample)
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def _scores_whole_pool(sel, ad, model, dev):
    """The selector's scoring pass for this process' pool, ignoring the process
    group (weak scaling: each rank's active dataset IS its own pool)."""
    import b200al.selectors as SM
    saved = SM.D.world
    SM.D.world = lambda: (0, 1)
    try:
        return sel._scores_local(ad, model, dev)
    finally:
        SM.D.world = saved


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
