"""Fixtures of bench.py / tools: synthetic pools, a seeded UNet state_dict in the reference's key layout and duck-typed
stand-ins for the reference's ActiveDataset / ExtendableDataset.  No arithmetic of the hot path lives here and
nothing is imported from oracle/ (the oracle is only the CPU arm / the checker).

Reference shapes: src/datasets/active_dataset.py:8-53, src/datasets/common.py:12-40 (dataset wrappers the selectors
consume), src/models/unet/unet.py:53-78,131-178 (module construction order = state_dict key order),
src/training/al_trainer.py:504-515 (the UNet al_train builds).
"""
from __future__ import annotations

import os
from collections import OrderedDict

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.abspath(__file__))
CHANNELS = (32, 64, 128, 256, 512)
SEED = 1337  # the reference CLI's default seed (src/entry/activelearning/train.py:11)


def case_name(i: int) -> str:
    return f"{i:06d}"


# ------------------------------------------------------------------------------------------------------- weights
def seeded_state_dict(num_out: int, seed: int = SEED) -> "OrderedDict[str, torch.Tensor]":
    """Default-initialised weights of the al_train UNet (2D, plain blocks, batch norm, transposed-conv upsampling)
    created in the reference's module order, with the BatchNorm tensors and the head of tests/golden/model.npz on top
    (running statistics calibrated on synthetic images, so activations are O(1) at every depth and the softmax is
    not saturated)."""
    torch.manual_seed(seed)
    sd: "OrderedDict[str, torch.Tensor]" = OrderedDict()

    def block(prefix, cin, cout, stride):
        conv = nn.Conv2d(cin, cout, 3, stride=stride, padding=1, bias=True)
        bn = nn.BatchNorm2d(cout)
        sd[f"{prefix}.0.weight"], sd[f"{prefix}.0.bias"] = conv.weight.detach(), conv.bias.detach()
        sd[f"{prefix}.2.weight"], sd[f"{prefix}.2.bias"] = bn.weight.detach(), bn.bias.detach()
        sd[f"{prefix}.2.running_mean"], sd[f"{prefix}.2.running_var"] = bn.running_mean.clone(), bn.running_var.clone()
        sd[f"{prefix}.2.num_batches_tracked"] = bn.num_batches_tracked.clone()

    prev = 1
    for l, c in enumerate(CHANNELS):
        block(f"encoder.levels.{l}.0.all", prev, c, 1 if l == 0 else 2)
        block(f"encoder.levels.{l}.1.all", c, c, 1)
        prev = c
    rev = CHANNELS[::-1]
    ups = []
    for l in range(4):
        ups.append(nn.ConvTranspose2d(rev[l], rev[l + 1], kernel_size=2, stride=2))
        block(f"decoder.levels.{l}.0.all", 2 * rev[l + 1], rev[l + 1], 1)
        block(f"decoder.levels.{l}.1.all", rev[l + 1], rev[l + 1], 1)
    for l, up in enumerate(ups):
        sd[f"decoder.upsamples.{l}.weight"], sd[f"decoder.upsamples.{l}.bias"] = up.weight.detach(), up.bias.detach()
    seg = nn.Conv2d(CHANNELS[0], num_out, 1)
    sd["decoder.seg_output.weight"], sd["decoder.seg_output.bias"] = seg.weight.detach(), seg.bias.detach()
    path = os.path.join(ROOT, "tests", "golden", "model.npz")
    if os.path.exists(path):
        g = np.load(path)
        with torch.no_grad():
            for k in sd:
                if "bn/" + k in g.files:
                    sd[k].copy_(torch.from_numpy(g["bn/" + k]))
            if f"head{num_out}/weight" in g.files:
                sd["decoder.seg_output.weight"].copy_(torch.from_numpy(g[f"head{num_out}/weight"]))
                sd["decoder.seg_output.bias"].copy_(torch.from_numpy(g[f"head{num_out}/bias"]))
    return sd


class StateDictModel(nn.Module):
    """What a selector needs from `model`: `state_dict()` in the reference's key layout, `eval()`, `modules()` for
    the hyper-parameters (BatchNorm eps, LeakyReLU slope, Dropout2d p as al_trainer.py:505-515 configures them).
    It has no forward: nothing on the bench's GPU arm can fall back to PyTorch."""

    def __init__(self, sd, dropout_p: float = 0.1):
        super().__init__()
        self._keys = list(sd.keys())
        for k, v in sd.items():
            self.register_buffer(k.replace(".", "__"), v.clone())
        self.hparams = nn.ModuleList([nn.BatchNorm2d(1, eps=1e-5), nn.LeakyReLU(0.01), nn.Dropout2d(dropout_p)])

    def state_dict(self, *a, **k):
        return OrderedDict((key, getattr(self, key.replace(".", "__"))) for key in self._keys)

    def forward(self, x):
        raise RuntimeError("StateDictModel has no forward: the acquisition pass runs in libb200al")


# --------------------------------------------------------------------------------------------------------- pools
def synthetic_image_u8(i: int, seed: int = SEED, size: int = 256) -> torch.Tensor:
    """[size,size] uint8: smooth low-frequency field + speckle with random contrast / brightness (8-bit like the
    PNG / JPG sources, fugc_dataset.py:140-143)."""
    g = torch.Generator().manual_seed(seed * 1_000_003 + i)
    z = torch.randn(1, 1, 16, 16, generator=g)
    field = F.interpolate(z, size=(size, size), mode="bilinear", align_corners=False)[0, 0]
    speckle = torch.randn(size, size, generator=g)
    cb = torch.rand(2, generator=g)
    img = (0.2 + 0.4 * cb[1]) + (0.3 + 0.7 * cb[0]) * 0.25 * (field + 0.35 * speckle)
    return (img.clamp(0, 1) * 255.0).round().to(torch.uint8)


def synthetic_pool_u8(n: int, start: int = 0, distinct: int = 256) -> torch.Tensor:
    """[n,256,256] uint8 on the CPU: `distinct` seeded base images, tiled with per-copy circular shifts so that every
    pool image is different (generating 100k independent images costs minutes of host time that no arm measures)."""
    base = torch.stack([synthetic_image_u8(start + i) for i in range(min(n, distinct))])
    reps = (n + base.shape[0] - 1) // base.shape[0]
    parts = [torch.roll(base, shifts=(7 * r, 13 * r), dims=(1, 2)) for r in range(reps)]
    return torch.cat(parts)[:n].contiguous()


class PoolBase(torch.utils.data.Dataset):
    """Stands in for FUGCDataset / BUSIDataset: `get_sample(i)` returns the reference's sample dict; `images_u8`
    ([N,H,W] uint8, host or device) is the resident fast path the drop-in selectors slice from."""

    def __init__(self, images_u8: torch.Tensor, names: list[str]):
        self.images_u8, self.names = images_u8, names

    def __len__(self):
        return len(self.names)

    def get_sample(self, index: int, normalize: bool = True):
        img = self.images_u8[index].to(torch.float32).div(255.0).unsqueeze(0)
        return {"image": img, "label": torch.zeros(img.shape[-2:], dtype=torch.long), "case_name": self.names[index]}

    def __getitem__(self, index):
        return self.get_sample(index)


class Extendable(torch.utils.data.Dataset):
    """ExtendableDataset (src/datasets/common.py:12-40): a list of case names over a base dataset."""

    def __init__(self, dataset: PoolBase, image_idx: list):
        self.dataset = dataset
        self.case_name_to_idx = {n: i for i, n in enumerate(dataset.names)}
        self.image_idx = image_idx

    def __len__(self):
        return len(self.image_idx)

    def get_sample(self, index: int, normalize: bool = True):
        return self.dataset.get_sample(self.case_name_to_idx[self.image_idx[index]], normalize)

    def __getitem__(self, index):
        return self.get_sample(index)


class Active:
    """ActiveDataset (src/datasets/active_dataset.py:8-53)."""

    def __init__(self, labeled_dataset: Extendable, pool_dataset: Extendable):
        self.labeled_dataset, self.pool_dataset = labeled_dataset, pool_dataset

    def extend_train_set(self, new_image_idx=()):
        self.labeled_dataset.image_idx.extend(new_image_idx)
        for idx in new_image_idx:
            self.pool_dataset.image_idx.remove(idx)

    def get_train_dataset(self):
        return self.labeled_dataset

    def get_pool_dataset(self):
        return self.pool_dataset

    def get_size(self):
        return len(self.labeled_dataset), len(self.pool_dataset)


def make_active(images_u8: torch.Tensor, n_labeled: int) -> Active:
    """The first `n_labeled` images are the labeled set, the rest the pool (positions keep dataset order)."""
    names = [case_name(i) for i in range(images_u8.shape[0])]
    base = PoolBase(images_u8, names)
    return Active(Extendable(base, names[:n_labeled]), Extendable(base, names[n_labeled:]))
