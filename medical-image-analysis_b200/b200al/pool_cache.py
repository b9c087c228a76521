"""Device-resident pool cache: the raw grayscale bytes of every pool image, uploaded once per run.

In the reference every acquisition round re-reads and re-resizes the whole pool on the host through a DataLoader
(`entropy_selector.py:30-39` -> `ExtendableDataset.get_sample` datasets/common.py:36-40 -> `FUGCDataset.get_sample`
fugc_dataset.py:122-164: PNG decode, `convert("L")`, `F.to_tensor`, `JointResize`, `ZScoreNormalize`).  The pool only
ever shrinks (`ActiveDataset.extend_train_set` active_dataset.py:41-44 moves case names out of it) and the pool
dataset is built without random transforms (`al_trainer.py:623`, include_transform=False), so its images are a pure
function of the files: decode them once, keep the uint8 pixels in HBM at their native (ragged) sizes, and run
resize + z-score statistics on the GPU per chunk (`b200al_prep_images`), with the normalisation itself applied in the
first conv's loader.

A cache is only built for datasets whose pixels can be obtained without the dataset's own float pipeline:
  * any base dataset exposing `b200_raw_u8(index) -> uint8 [H,W] array` (plus `image_size`, `normalize`), or
  * the reference's `FUGCDataset` / `BUSIDataset` (same file naming as their `get_sample`),
and only when `transform` is None, `image_channels == 1`, `image_size` is set and `normalize` is None or the
reference's `ZScoreNormalize`.  Everything else keeps using the DataLoader path.
"""
from __future__ import annotations

import logging
import time

import numpy as np
import torch

log = logging.getLogger("b200al")


def _hw(image_size) -> tuple[int, int] | None:
    if image_size is None:
        return None
    if isinstance(image_size, int):
        return (image_size, image_size)
    s = list(image_size)
    if len(s) == 1:
        s = s * 2
    return (int(s[0]), int(s[1]))


def _raw_reader(base):
    """index -> uint8 [H,W] numpy array, or None when the dataset's files cannot be read directly."""
    fn = getattr(base, "b200_raw_u8", None)
    if callable(fn):
        return lambda i: np.ascontiguousarray(fn(i), dtype=np.uint8)
    kind = type(base).__name__
    if kind not in ("FUGCDataset", "BUSIDataset") or not hasattr(base, "data_path") or not hasattr(base, "samples_list"):
        return None
    from PIL import Image
    cls = type(base)
    if kind == "FUGCDataset":  # fugc_dataset.py:125-139
        split_dir = {"train": cls.TRAIN_DIR, "valid": cls.VALID_DIR}.get(base.split, cls.TEST_DIR)
        folder = base.data_path / split_dir / cls.IMAGES_DIR
    else:  # busi_dataset.py:123
        folder = base.data_path / cls.IMAGES_DIR

    def read(i):
        with Image.open(folder / f"{base.samples_list[i]}.png") as im:
            return np.asarray(im.convert("L"), dtype=np.uint8)

    return read


class DevicePoolCache:
    def __init__(self, raws: list[np.ndarray], out_hw: tuple[int, int], normalize: bool, device):
        self.device = torch.device(device)
        self.out_hw, self.normalize = out_hw, bool(normalize)
        sizes = np.array([r.shape for r in raws], dtype=np.int64).reshape(-1, 2)
        nbytes = sizes[:, 0] * sizes[:, 1]
        offs = np.zeros(len(raws), dtype=np.int64)
        offs[1:] = np.cumsum(nbytes)[:-1]
        host = torch.empty(int(nbytes.sum()), dtype=torch.uint8)
        flat = host.numpy()
        for r, o, n in zip(raws, offs, nbytes):
            flat[o:o + n] = r.reshape(-1)
        self.raw = host.to(self.device)
        self.offsets = torch.from_numpy(offs).to(self.device)
        self.hw = torch.from_numpy(sizes.astype(np.int32)).to(self.device)
        self.max_h, self.max_w = (int(sizes[:, 0].max()), int(sizes[:, 1].max())) if len(raws) else (1, 1)
        self.n, self.nbytes = len(raws), int(nbytes.sum())

    @classmethod
    def try_build(cls, base, device) -> "DevicePoolCache | None":
        if getattr(base, "transform", None) is not None or getattr(base, "image_channels", 1) != 1:
            return None
        out_hw = _hw(getattr(base, "image_size", None))
        norm = getattr(base, "normalize", None)
        if out_hw is None or (norm is not None and type(norm).__name__ != "ZScoreNormalize"):
            return None
        if norm is not None and getattr(norm, "target_dtype", torch.float32) != torch.float32:
            return None
        read = _raw_reader(base)
        if read is None:
            return None
        t0 = time.perf_counter()
        raws = [read(i) for i in range(len(base))]
        cache = cls(raws, out_hw, norm is not None, device)
        log.info("pool cache: %d images, %.1f MB raw uint8 resident on %s, built in %.2f s", cache.n,
                 cache.nbytes / 1e6, cache.device, time.perf_counter() - t0)
        return cache

    def gather(self, eng, rows: torch.Tensor):
        """Model-ready images of base-dataset rows (int64 tensor): (fp32 [m,H,W] on device, norm [m,2] or None)."""
        rows = rows.to(self.device)
        return eng.prep(self.raw, self.offsets.index_select(0, rows).contiguous(),
                        self.hw.index_select(0, rows).contiguous(), self.max_h, self.max_w, self.normalize)


def pool_cache_for(base, device, mode: str = "auto") -> DevicePoolCache | None:
    """The cache attached to a base dataset, building it on first use ("auto"); None when unsupported or "off"."""
    if mode == "off" or base is None:
        return None
    dev = torch.device(device)
    if dev.type == "cuda" and dev.index is None:
        dev = torch.device("cuda", torch.cuda.current_device())
    cache = getattr(base, "_b200_pool_cache", None)
    if cache is not None and cache.device == dev and cache.n == len(base):
        return cache
    cache = DevicePoolCache.try_build(base, dev)
    if cache is not None:
        try:
            base._b200_pool_cache = cache
        except AttributeError:
            pass
    return cache
