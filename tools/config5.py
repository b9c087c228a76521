"""BASELINE.json configs[4]: a full al_train-style run - `num_rounds` rounds of (acquisition with the drop-in selector,
then UNet training) on synthetic BUSI-shape data written as PNG files - reporting acquisition vs training seconds per
round.  Mirrors the reference's round loop (src/training/al_trainer.py:1041-1115):

    on_round_start: [round > 0: load last round's weights] -> select_next_batch(active_dataset, budget, model, device)
                    -> extend_train_set -> [round > 0: rebuild the model] -> train

Round 0 has no labeled data, so the selector takes its cold-start branch (torch.rand on the seeded CPU generator,
entropy_selector.py:64-71), exactly like the reference.  The reference tree does not exist on the GPU box, so the
trainer side is STOCK PYTORCH here (a plain nn.Module UNet with the reference's state_dict layout, Adam, CE + soft
Dice, cuDNN convs) - it is the caller of the hot path, not part of it; only acquisition runs in libb200al.
Multi-GPU (torchrun): rank 0 trains, every rank joins the acquisition (the selector broadcasts rank 0's weights and
shards the pool); the others wait at the barrier during training ("replicas only" for training).

  python tools/config5.py [--pool 2048] [--budget 32] [--rounds 3] [--iters 60] [--selector b200-entropy]
  torchrun --nproc-per-node 8 tools/config5.py ...
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import tempfile
import time

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "medical-image-analysis_b200")]

CH = (32, 64, 128, 256, 512)


class PlainBlock(nn.Module):  # blocks.py:66-105: conv -> Dropout2d -> BatchNorm2d -> LeakyReLU as `all`
    def __init__(self, cin, cout, stride, p):
        super().__init__()
        self.all = nn.Sequential(nn.Conv2d(cin, cout, 3, stride, 1, bias=True), nn.Dropout2d(p), nn.BatchNorm2d(cout),
                                 nn.LeakyReLU(0.01, inplace=True))

    def forward(self, x):
        return self.all(x)


class Encoder(nn.Module):
    def __init__(self, p):
        super().__init__()
        self.levels = nn.ModuleList()
        prev = 1
        for l, c in enumerate(CH):
            self.levels.append(nn.Sequential(PlainBlock(prev, c, 1 if l == 0 else 2, p), PlainBlock(c, c, 1, p)))
            prev = c


class Decoder(nn.Module):
    def __init__(self, K, p):
        super().__init__()
        rev = CH[::-1]
        self.levels, self.upsamples = nn.ModuleList(), nn.ModuleList()
        for l in range(4):
            self.upsamples.append(nn.ConvTranspose2d(rev[l], rev[l + 1], 2, 2))
            self.levels.append(nn.Sequential(PlainBlock(2 * rev[l + 1], rev[l + 1], 1, p), PlainBlock(rev[l + 1], rev[l + 1], 1, p)))
        self.seg_output = nn.Conv2d(CH[0], K, 1)


class TorchUNet(nn.Module):
    """Stock-PyTorch UNet with the reference's module / state_dict layout (unet.py:53-78, 131-178, 208-224)."""

    def __init__(self, K=2, p=0.1):
        super().__init__()
        self.encoder, self.decoder = Encoder(p), Decoder(K, p)

    def forward(self, x):
        skips = []
        for s in self.encoder.levels:
            x = s(x)
            skips.append(x)
        for l in range(4):
            x = self.decoder.upsamples[l](x)
            x = self.decoder.levels[l](torch.cat([skips[3 - l], x], 1))
        return self.decoder.seg_output(x)


class PngDataset(torch.utils.data.Dataset):
    """BUSI-like folder of grayscale PNGs + masks (busi_dataset.py:94-143).  `b200_raw_u8` hands the decoded 8-bit
    pixels to the device pool cache (b200al/pool_cache.py); `get_sample` is the float pipeline the TRAINER uses."""

    def __init__(self, folder, names, size=256):
        self.folder, self.names, self.image_size = folder, names, size
        self.transform, self.normalize, self.image_channels = None, None, 1

    def __len__(self):
        return len(self.names)

    def b200_raw_u8(self, index):
        import cv2
        return cv2.imread(os.path.join(self.folder, f"{self.names[index]}.png"), cv2.IMREAD_GRAYSCALE)

    def get_sample(self, index, normalize=True):
        import cv2
        img = torch.from_numpy(self.b200_raw_u8(index)).float().div(255)[None, None]
        lab = torch.from_numpy(cv2.imread(os.path.join(self.folder, f"{self.names[index]}_mask.png"), cv2.IMREAD_GRAYSCALE))
        s = self.image_size
        img = F.interpolate(img, size=(s, s), mode="bilinear", antialias=True, align_corners=False)[0]
        lab = F.interpolate(lab[None, None].float(), size=(s, s), mode="nearest")[0, 0].long().clamp_(0, 1)
        return {"image": img, "label": lab, "case_name": self.names[index]}

    def __getitem__(self, index):
        return self.get_sample(index)


def write_dataset(folder, n, seed=1337):
    """n synthetic BUSI-shaped cases (ragged sizes around 500 x 600 like the real set) as PNG + blob mask."""
    import cv2
    from bench_fixtures import synthetic_image_u8
    rng = np.random.default_rng(seed)
    names = []
    for i in range(n):
        h, w = int(rng.integers(380, 620)), int(rng.integers(440, 720))
        base = synthetic_image_u8(i, seed=seed).numpy()
        img = cv2.resize(base, (w, h), interpolation=cv2.INTER_LINEAR)
        yy, xx = np.mgrid[0:h, 0:w]
        cy, cx, r = rng.uniform(0.3, 0.7) * h, rng.uniform(0.3, 0.7) * w, rng.uniform(0.08, 0.2) * min(h, w)
        mask = (((yy - cy) ** 2 + (xx - cx) ** 2) < r * r).astype(np.uint8)
        img = np.where(mask > 0, (img * 0.55).astype(np.uint8), img)   # a darker lesion the net can learn
        name = f"case_{i:05d}"
        cv2.imwrite(os.path.join(folder, f"{name}.png"), img)
        cv2.imwrite(os.path.join(folder, f"{name}_mask.png"), mask * 255)
        names.append(name)
    return names


def soft_dice(logits, target, smooth=1e-5):
    prob = logits.softmax(1)
    onehot = F.one_hot(target, prob.shape[1]).permute(0, 3, 1, 2).float()
    inter = (prob * onehot).sum((2, 3))
    return (1 - (2 * inter + smooth) / (prob.sum((2, 3)) + onehot.sum((2, 3)) + smooth)).mean()


def train(model, dataset, names, iters, batch, dev, seed):
    g = torch.Generator().manual_seed(seed)
    opt = torch.optim.Adam(model.parameters(), lr=1e-3)
    model.train()
    idx_of = {n: i for i, n in enumerate(dataset.names)}
    loss = torch.zeros(())
    for it in range(iters):
        pick = torch.randint(0, len(names), (min(batch, len(names)),), generator=g).tolist()
        samples = [dataset.get_sample(idx_of[names[j]]) for j in pick]
        x = torch.stack([s["image"] for s in samples]).to(dev)
        y = torch.stack([s["label"] for s in samples]).to(dev)
        out = model(x)
        loss = F.cross_entropy(out, y) + soft_dice(out, y)
        opt.zero_grad(set_to_none=True)
        loss.backward()
        opt.step()
    return float(loss)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--pool", type=int, default=2048)
    ap.add_argument("--budget", type=int, default=32)
    ap.add_argument("--rounds", type=int, default=3)
    ap.add_argument("--iters", type=int, default=60)
    ap.add_argument("--batch", type=int, default=12)   # the reference CLI's default batch size
    ap.add_argument("--selector", default="b200-entropy")
    ap.add_argument("--precision", default="mixed")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "r02_config5.jsonl"))
    args = ap.parse_args()
    import torch.distributed as dist
    import b200al
    from bench_fixtures import Active, Extendable
    world, rank = int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("RANK", 0))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    torch.manual_seed(1337)   # al_trainer.py:412-419 seeds the CPU generator the cold start draws from
    np.random.seed(1337)
    folder = os.path.join(tempfile.gettempdir(), f"b200al_config5_{args.pool}")
    if rank == 0 and not os.path.exists(os.path.join(folder, "names.json")):
        os.makedirs(folder, exist_ok=True)
        json.dump(write_dataset(folder, args.pool), open(os.path.join(folder, "names.json"), "w"))
    if world > 1:
        dist.barrier()
    names = json.load(open(os.path.join(folder, "names.json")))
    base = PngDataset(folder, names)
    active = Active(Extendable(base, []), Extendable(base, list(names)))
    kw = dict(batch_size=args.batch, num_workers=0, pin_memory=True, chunk_samples=128, precision=args.precision)
    selector = b200al.SELECTORS[args.selector](**kw)
    model = TorchUNet(2).to(dev)
    records = []
    for rnd in range(args.rounds):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        picked = selector.select_next_batch(active, args.budget, model, dev)     # al_trainer.py:1061-1066
        torch.cuda.synchronize()
        t_acq = time.perf_counter() - t0
        labeled_before, pool_before = active.get_size()
        active.extend_train_set(picked)                                         # :1068
        if rnd > 0:
            torch.manual_seed(1337 + rnd)
            model = TorchUNet(2).to(dev)                                        # :1075 rebuild (no persist_model_weight)
        t0 = time.perf_counter()
        loss = train(model, base, active.get_train_dataset().image_idx, args.iters, args.batch, dev, rnd) if rank == 0 else 0.0
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        t_train = time.perf_counter() - t0
        rec = {"round": rnd, "selector": args.selector, "precision": args.precision, "n_gpus": world,
               "pool": pool_before, "labeled": labeled_before, "budget": args.budget, "picked": len(picked),
               "cold_start": labeled_before == 0, "acquisition_s": round(t_acq, 4), "training_s": round(t_train, 3),
               "train_iters": args.iters, "final_loss": round(loss, 4),
               "acquisition_images_per_s": round(pool_before / t_acq, 1),
               "pool_cache": "device-resident raw uint8 cache (decoded once in round 1, reused afterwards)"}
        records.append(rec)
        if rank == 0:
            print(json.dumps(rec), flush=True)
    if rank == 0:
        os.makedirs(os.path.dirname(args.out), exist_ok=True)
        with open(args.out, "a") as f:
            for r in records:
                f.write(json.dumps(r) + "\n")
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
