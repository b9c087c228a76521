"""Seeded synthetic FUGC-/BUSI-shaped pools (oracle side; test-only).

Shapes follow the reference datasets after `JointResize(256)`
(src/entry/activelearning/train.py:30; src/datasets/fugc/fugc_dataset.py:85-89,
140-164): grayscale [1,256,256] fp32 in [0,1] coming from 8-bit images
(`to_tensor` == uint8/255), label [256,256] int64, case_name str.

Image i depends only on (seed, i), so any pool shard regenerates its slice.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn.functional as F
from torch.utils.data import Dataset

SEED = 1337  # reference CLI default seed (src/entry/activelearning/train.py:11)
H = W = 256


def synth_image_u8(i: int, seed: int = SEED, size: int = H) -> torch.Tensor:
    """[size,size] uint8: smooth low-frequency field + speckle, random contrast."""
    g = torch.Generator().manual_seed(seed * 1_000_003 + i)
    z = torch.randn(1, 1, 16, 16, generator=g)
    field = F.interpolate(z, size=(size, size), mode="bilinear", align_corners=False)[0, 0]
    speckle = torch.randn(size, size, generator=g)
    cb = torch.rand(2, generator=g)
    contrast = 0.3 + 0.7 * cb[0]
    bright = 0.2 + 0.4 * cb[1]
    img = bright + contrast * 0.25 * (field + 0.35 * speckle)
    return (img.clamp(0, 1) * 255.0).round().to(torch.uint8)


def synth_pool_u8(n: int, start: int = 0, seed: int = SEED, size: int = H) -> torch.Tensor:
    """[n,size,size] uint8 pool slice [start, start+n)."""
    return torch.stack([synth_image_u8(start + i, seed, size) for i in range(n)])


def case_name(i: int) -> str:
    return f"{i:06d}"


class SynthBase(Dataset):
    """Stands in for FUGCDataset/BUSIDataset: `get_sample(i)` returns the
    reference's sample dict {"image","label","case_name"}."""

    def __init__(self, images_u8: torch.Tensor, names: list[str] | None = None):
        self.images_u8 = images_u8
        self.names = names if names is not None else [case_name(i) for i in range(len(images_u8))]

    @staticmethod
    def find_samples(data_path, require_label=True):
        raise RuntimeError("synthetic dataset")

    def __len__(self):
        return len(self.names)

    def get_sample(self, index: int, normalize: bool = True):
        img = self.images_u8[index].to(torch.float32).div(255.0).unsqueeze(0)
        return {
            "image": img,
            "label": torch.zeros(img.shape[-2:], dtype=torch.long),
            "case_name": self.names[index],
        }

    def __getitem__(self, index):
        return self.get_sample(index)


class ZScoreNormalize:
    """Marker mirror of the reference's transform (src/transforms/normalization.py:9); SynthRawBase applies the
    oracle's restatement of it."""
    target_dtype = torch.float32


class SynthRawBase(Dataset):
    """Stands in for FUGCDataset/BUSIDataset built the way `al_train` builds the POOL dataset
    (al_trainer.py:623: transform=None, image_size=--image-size, normalize=ZScoreNormalize() under --do-normalize):
    ragged grayscale uint8 "files"; `get_sample` runs the oracle's restatement of to_tensor -> JointResize ->
    ZScoreNormalize on the CPU (oracle/prep_ref.py).  `b200_raw_u8` is the raw-pixel hook the device pool cache
    reads instead."""

    def __init__(self, raws: list, image_size=256, normalize: bool = False, names: list[str] | None = None):
        self.raws = raws
        self.names = names if names is not None else [case_name(i) for i in range(len(raws))]
        self.image_size = image_size
        self.normalize = ZScoreNormalize() if normalize else None
        self.transform = None
        self.image_channels = 1

    def __len__(self):
        return len(self.names)

    def b200_raw_u8(self, index: int):
        return self.raws[index]

    def get_sample(self, index: int, normalize: bool = True):
        from .prep_ref import prep_image
        hw = (self.image_size, self.image_size) if isinstance(self.image_size, int) else tuple(self.image_size)
        img = prep_image(self.raws[index], hw, self.normalize is not None and normalize)
        return {
            "image": torch.from_numpy(img).unsqueeze(0),
            "label": torch.zeros(hw, dtype=torch.long),
            "case_name": self.names[index],
        }

    def __getitem__(self, index):
        return self.get_sample(index)


class MirrorExtendable(Dataset):
    """Behavioural mirror of ExtendableDataset (src/datasets/common.py:12-40):
    a list of case names over a base dataset (without the reference's habit of
    loading every sample once just to read its name, common.py:21-24)."""

    def __init__(self, dataset: SynthBase, image_idx: list | None = None):
        self.dataset = dataset
        self.case_name_to_idx = {n: i for i, n in enumerate(dataset.names)}
        self.image_idx = list(self.case_name_to_idx.keys()) if image_idx is None else image_idx

    def __len__(self):
        return len(self.image_idx)

    def __getitem__(self, index):
        return self.get_sample(index)

    def get_sample(self, index: int, normalize: bool = True):
        return self.dataset.get_sample(self.case_name_to_idx[self.image_idx[index]], normalize)


class MirrorActive:
    """Behavioural mirror of ActiveDataset (src/datasets/active_dataset.py:8-53)."""

    def __init__(self, labeled_dataset, pool_dataset):
        self.labeled_dataset = labeled_dataset
        self.pool_dataset = pool_dataset

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


def make_active(images_u8: torch.Tensor, n_labeled: int, reference=None):
    """Pool split the way al_train starts a round: the first `n_labeled` names
    are labeled, the rest form the pool (positions keep dataset order).
    With `reference` (oracle.ref_import.load_reference()) the reference's own
    ExtendableDataset/ActiveDataset types are used instead of the mirrors."""
    base = SynthBase(images_u8)
    names = list(base.names)
    if reference is not None:
        class _RefBase(SynthBase, reference.BaseDataset):  # a BaseDataset to the reference
            pass
        base = _RefBase(images_u8)
        lab = reference.ExtendableDataset(base, names[:n_labeled])
        pool = reference.ExtendableDataset(base, names[n_labeled:])
        return reference.ActiveDataset(lab, pool)
    return MirrorActive(MirrorExtendable(base, names[:n_labeled]), MirrorExtendable(base, names[n_labeled:]))


# ---------------------------------------------------------------------------
# MC-dropout keep masks: counter-based hash shared with the CUDA kernel
# (medical-image-analysis_b200/csrc/common.cuh: mc_keep_hash).  Position = pool
# position of the image, t = repeat, ch = global dropout channel (0..2943).
# ---------------------------------------------------------------------------
_M1 = np.uint64(0x9E3779B97F4A7C15)
_M2 = np.uint64(0xBF58476D1CE4E5B9)
_M3 = np.uint64(0x94D049BB133111EB)


def mc_hash24(seed: int, pos: np.ndarray, t: np.ndarray, ch: np.ndarray) -> np.ndarray:
    with np.errstate(over="ignore"):
        pos = pos.astype(np.uint64)
        tc = (t.astype(np.uint64) << np.uint64(32)) | ch.astype(np.uint64)
        x = (np.uint64(seed) * _M1) ^ (pos * _M2) ^ (tc * _M3)
        x ^= x >> np.uint64(30)
        x *= _M2
        x ^= x >> np.uint64(27)
        x *= _M3
        x ^= x >> np.uint64(31)
        return (x >> np.uint64(40)).astype(np.uint32)


def mc_threshold(p: float) -> int:
    return int(np.floor(float(p) * (1 << 24)))


def mc_masks(seed: int, positions, T: int, p: float, n_ch: int = 2944) -> torch.Tensor:
    """[len(positions), T, n_ch] fp32 multipliers: 0 or 1/(1-p) (Dropout2d's
    train-mode scaling), keep iff hash24 >= floor(p*2^24)."""
    pos = np.asarray(positions, dtype=np.uint64)[:, None, None]
    t = np.arange(T, dtype=np.uint64)[None, :, None]
    ch = np.arange(n_ch, dtype=np.uint64)[None, None, :]
    keep = mc_hash24(seed, pos + 0 * t + 0 * ch, t + 0 * pos + 0 * ch, ch + 0 * pos + 0 * t) >= mc_threshold(p)
    scale = np.float32(1.0) / (np.float32(1.0) - np.float32(p))
    return torch.from_numpy(keep.astype(np.float32) * scale)
