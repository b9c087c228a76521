"""CPU restatement of the reference's pool-image input pipeline (oracle side; TEST INFRASTRUCTURE ONLY -
nothing in the product path may import this).

Path restated (reference file:line):
  * `FUGCDataset.get_sample` / `BUSIDataset.get_sample`  (src/datasets/fugc/fugc_dataset.py:140-164,
    src/datasets/busi/busi_dataset.py:120-148): PNG -> `convert("L")` -> `F.to_tensor` (uint8/255, fp32)
    -> `JointResize(image_size)` -> optional `ZScoreNormalize`.
  * `JointResize.__call__` (src/transforms/joint_transform.py:20-25): `F.resize(image, size, BILINEAR)` on a float
    TENSOR, i.e. torchvision (unpinned, `pyproject.toml:19`; 0.26 here) -> `torch.nn.functional.interpolate(
    mode="bilinear", align_corners=False, antialias=True)` -> ATen `_upsample_bilinear2d_aa`.  That algorithm is
    third-party (PyTorch ATen, aten/src/ATen/native/cpu/UpSampleKernel.cpp, `_compute_indices_min_size_weights_aa`
    + separable horizontal-then-vertical passes); restated here from its published definition:
        scale   = in / out                      support = scale if scale >= 1 else 1
        center  = scale * (i + 0.5)             invscale = 1/scale if scale >= 1 else 1
        xmin    = max(0, int(center - support + 0.5))
        xsize   = min(in, int(center + support + 0.5)) - xmin
        w_j     = tri((j + xmin - center + 0.5) * invscale),  tri(x) = max(0, 1 - |x|),  normalised to sum 1
  * `ZScoreNormalize.__call__` (src/transforms/normalization.py:13-21): `(image - image.mean()) /
    image.std().clip(1e-8)` with torch's default UNBIASED std over all C*H*W elements.

Pinned by tests/test_oracle_vs_reference.py against torch's interpolate, torchvision's `F.resize` and the reference's
own `JointResize` / `ZScoreNormalize` classes, and by tests/golden/prep.npz (made by oracle/make_golden_prep.py from
the reference classes) on machines without /root/reference.
"""
from __future__ import annotations

import numpy as np


def aa_taps(in_size: int, out_size: int):
    """Per output index: (xmin [out] int64, xsize [out] int64, weights [out, max_taps] float32)."""
    f32 = np.float32
    scale = f32(in_size) / f32(out_size)
    support = scale if scale >= 1.0 else f32(1.0)
    invscale = f32(1.0) / scale if scale >= 1.0 else f32(1.0)
    max_taps = int(np.ceil(support)) * 2 + 1
    xmin = np.zeros(out_size, np.int64)
    xsize = np.zeros(out_size, np.int64)
    w = np.zeros((out_size, max_taps), f32)
    for i in range(out_size):
        center = scale * f32(i + 0.5)
        lo = max(int(center - support + f32(0.5)), 0)
        n = min(int(center + support + f32(0.5)), in_size) - lo
        n = min(max(n, 0), max_taps)
        j = np.arange(n, dtype=f32)
        wj = np.maximum(f32(0.0), f32(1.0) - np.abs((j + f32(lo) - center + f32(0.5)) * invscale)).astype(f32)
        tot = f32(0.0)
        for v in wj:  # sequential fp32 sum like ATen
            tot = f32(tot + v)
        if tot != 0:
            wj = (wj * (f32(1.0) / tot)).astype(f32)
        xmin[i], xsize[i] = lo, n
        w[i, :n] = wj
    return xmin, xsize, w


def resize_bilinear_aa(img: np.ndarray, out_hw: tuple[int, int]) -> np.ndarray:
    """img float32 [Hs, Ws] -> float32 [Ho, Wo]; horizontal pass, then vertical, fp32 throughout."""
    img = np.asarray(img, np.float32)
    Hs, Ws = img.shape
    Ho, Wo = out_hw
    if (Hs, Ws) == (Ho, Wo):
        return img.copy()
    cur = img
    if Ws != Wo:
        xmin, xsize, w = aa_taps(Ws, Wo)
        tmp = np.zeros((Hs, Wo), np.float32)
        for j in range(w.shape[1]):
            idx = np.minimum(xmin + j, Ws - 1)
            tmp += (cur[:, idx] * np.where(j < xsize, w[:, j], np.float32(0))[None, :]).astype(np.float32)
        cur = tmp
    if Hs != Ho:
        ymin, ysize, w = aa_taps(Hs, Ho)
        out = np.zeros((Ho, cur.shape[1]), np.float32)
        for j in range(w.shape[1]):
            idx = np.minimum(ymin + j, Hs - 1)
            out += (cur[idx, :] * np.where(j < ysize, w[:, j], np.float32(0))[:, None]).astype(np.float32)
        cur = out
    return cur


def zscore(img: np.ndarray) -> np.ndarray:
    """(x - mean) / clip(std_unbiased, 1e-8); statistics in fp64, arithmetic on the image in fp32."""
    x = np.asarray(img, np.float32)
    mean = np.float32(x.astype(np.float64).mean())
    std = np.float32(x.astype(np.float64).std(ddof=1)) if x.size > 1 else np.float32(np.nan)
    return ((x - mean) / np.maximum(std, np.float32(1e-8))).astype(np.float32)


def prep_image(raw_u8: np.ndarray, out_hw: tuple[int, int], normalize: bool) -> np.ndarray:
    """Grayscale uint8 [Hs, Ws] -> the [Ho, Wo] float32 image the model sees (channel dim left out)."""
    x = (np.asarray(raw_u8, np.uint8).astype(np.float32) / np.float32(255.0)).astype(np.float32)  # F.to_tensor
    x = resize_bilinear_aa(x, out_hw)
    return zscore(x) if normalize else x


def synth_raw_u8(i: int, seed: int = 1337) -> np.ndarray:
    """Seeded ragged grayscale image: FUGC-like 336x544, BUSI-like ~500x580 with per-image jitter, and a few small
    ones that need UP-sampling.  Depends only on (seed, i)."""
    rng = np.random.default_rng(seed * 7_919 + i)
    kind = i % 4
    if kind == 0:
        h, w = 336, 544
    elif kind == 1:
        h, w = int(rng.integers(430, 620)), int(rng.integers(450, 760))
    elif kind == 2:
        h, w = int(rng.integers(150, 250)), int(rng.integers(180, 256))
    else:
        h, w = int(rng.integers(257, 330)), int(rng.integers(200, 300))
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    f = rng.uniform(0.01, 0.06, 4).astype(np.float32)
    field = np.sin(yy * f[0] + xx * f[1]) + np.cos(yy * f[2] - xx * f[3])
    speck = rng.standard_normal((h, w)).astype(np.float32)
    img = 0.45 + 0.18 * field + 0.12 * speck
    return (np.clip(img, 0, 1) * 255.0).round().astype(np.uint8)
